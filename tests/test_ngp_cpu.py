"""CPU-only checks of the Instant-NGP variant (scripts/instantngp.py): the host-side level table of the C ABI against the
oracle's, properties of the oracle itself (PARITY UNPINNED: tiny-cuda-nn is absent, so the oracle is held to the
published algorithm's invariants instead of golden vectors), and the module's constructor / state-dict contract."""
import numpy as np
import pytest
import torch

from oracle import ngp_oracle as N


@pytest.fixture(scope="module")
def D():
    import __graft_entry__ as ge
    ge.build()
    import deepfields_nerf_b200 as d
    return d


@pytest.mark.parametrize("log2", [15, 17, 19])
def test_level_table_matches_oracle(D, log2):
    g = D.ops.hashgrid_levels(log2)
    scale, res, size, offset, total = N.level_table(log2)
    assert np.array_equal(np.array(g.scale, np.float32), scale)
    assert list(g.res) == list(res) and list(g.size) == list(size) and list(g.offset) == list(offset)
    assert g.n_entries == total and g.log2_hashmap_size == log2
    assert list(res[:5]) == [16, 24, 36, 54, 81] and res[15] == 7007   # 16 * 1.5^l, SURVEY 8c
    assert all(s % 8 == 0 for s in size) and max(size) == 1 << log2


def test_level_table_rejects_bad_size(D):
    with pytest.raises(ValueError):
        D.ops.hashgrid_levels(40)


def test_hash_known_answers():
    # coherent_prime_hash with primes {1, 2654435761, 805459861}: computed by hand in Python integers
    c = np.array([[1, 1, 1], [5, 7, 11], [4000, 123, 77]], np.uint32)
    want = [(int(x) ^ (int(y) * 2654435761 & 0xFFFFFFFF) ^ (int(z) * 805459861 & 0xFFFFFFFF)) % (1 << 19) for x, y, z in c]
    assert list(N._grid_index(c, 7007, 1 << 19)) == want
    # a dense level indexes x + y*res + z*res^2
    assert list(N._grid_index(c[:2], 16, 4096)) == [1 + 16 + 256, 5 + 7 * 16 + 11 * 256]


def test_encode_invariants():
    rng = np.random.default_rng(0)
    total = N.level_table(15)[4]
    pos = rng.uniform(0, 1, (257, 3)).astype(np.float32)
    # partition of unity: a constant table interpolates to the constant
    const = np.full((total, 2), 0.25, np.float16)
    assert np.abs(N.hashgrid_encode(pos, const, 15).astype(np.float32) - 0.25).max() <= 2e-4
    # a point on a vertex of dense level 0 (scale 15: x*15 + 0.5 => vertex k sits at (k - 0.5) / 15) reads that entry
    tab = rng.uniform(-1, 1, (total, 2)).astype(np.float16)
    k = np.array([[3, 5, 9]], np.float32)
    e = N.hashgrid_encode(((k - 0.5) / 15).astype(np.float32), tab, 15)
    assert np.allclose(e[0, :2].astype(np.float32), tab[3 + 5 * 16 + 9 * 256].astype(np.float32), atol=2e-3)


def test_backward_is_the_adjoint_of_encode():
    rng = np.random.default_rng(1)
    total = N.level_table(15)[4]
    pos = rng.uniform(0, 1, (300, 3)).astype(np.float32)
    tab = rng.uniform(-1, 1, (total, 2)).astype(np.float16)
    d = rng.uniform(-1, 1, (300, 32)).astype(np.float16)
    lhs = float((N.hashgrid_encode(pos, tab, 15).astype(np.float64) * d.astype(np.float64)).sum())
    rhs = float((N.hashgrid_backward(pos, d, 15) * tab.astype(np.float64)).sum())
    assert abs(lhs - rhs) <= 2e-3 * max(1.0, abs(lhs))      # fp16 rounding of the encode output only


def test_sh4_known_values():
    # direction +z (input 0.5, 0.5, 1.0): Y00, Y10 = 0.4886 z, Y20 = 0.9462 z^2 - 0.3154, Y30 = 0.3732 z (5 z^2 - 3)
    o = N.sh4(np.array([[0.5, 0.5, 1.0]], np.float32)).astype(np.float32)[0]
    want = np.zeros(16, np.float32)
    want[0], want[2], want[6], want[12] = 0.28209479, 0.48860251, 0.63078313, 0.74635267
    assert np.allclose(o, want, atol=1e-3)


def test_module_contract(D):
    with pytest.raises(ValueError):          # instantngp.py:20-21
        D.InstantNGP("huge")
    m = D.InstantNGP("small")
    sd = m.state_dict()
    assert list(sd.keys()) == ["pos_encoding.params", "dir_encoding.params", "dmlp.params", "cmlp.params"]
    assert sd["pos_encoding.params"].numel() == 2 * N.level_table(15)[4] and sd["dir_encoding.params"].numel() == 0
    assert sd["dmlp.params"].numel() == 3072 and sd["cmlp.params"].numel() == 7168
    assert float(sd["pos_encoding.params"].abs().max()) <= 1e-4
    assert m.pos_encoding.n_output_dims == 32 and m.dir_encoding.n_output_dims == 16
    with pytest.raises(RuntimeError):        # no CPU fallback
        m(torch.zeros(4, 3), torch.zeros(4, 3))
