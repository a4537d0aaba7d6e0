"""CPU-only checks: the C-ABI library loads and exports every symbol include/deepfields_b200.h declares, the
host-side mirrors keep the reference's names / state-dict layout, the config loader resolves ${} references.
No compute call is made (there is no GPU here)."""
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def D():
    import __graft_entry__ as ge
    ge.build()
    import deepfields_nerf_b200 as d
    return d


def test_header_symbols_exported(D):
    hdr = open(os.path.join(ROOT, "include", "deepfields_b200.h")).read()
    declared = set(re.findall(r"\b(dfn_[a-z0-9_]+)\s*\(", hdr))
    declared.discard("dfn_point_source")
    assert len(declared) >= 24
    lib = D._lib.load()
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
        assert name in D._lib.SIGNATURES, f"{name} has no ctypes signature"
    assert lib.dfn_abi_version() == 1
    assert lib.dfn_mlp_fwd_pack_bytes() == 1265664
    assert lib.dfn_mlp_stash_bytes(129) == 2 * (9 * 65536 + 32768 + 8 * 4096)  # 2 tiles: h1..h8, features, colour, bit masks


def test_state_dict_layout_matches_reference(D, golden):
    from common import make_state_dict
    sd = make_state_dict(1, 256)
    m = D.OptimizedVanillaNeRF()
    assert list(m.state_dict().keys()) == list(sd.keys())
    for k, v in m.state_dict().items():
        assert v.shape == sd[k].shape and v.dtype == torch.float32, k
    m.load_state_dict(sd)
    flat = m.flatten_parameters()
    assert flat.numel() == 595844 == D.ops.MLP_NUM_PARAMS
    assert torch.equal(flat[: 256 * 63].view(256, 63), sd["pos_layers.0.weight"])
    m.pos_layers[0].weight.data.add_(1.0)             # parameters alias the flat buffer
    assert torch.equal(flat[: 256 * 63].view(256, 63), sd["pos_layers.0.weight"] + 1.0)
    assert float(m.density_layer.bias.detach()) == pytest.approx(0.1)  # after load: golden init value
    c, f = D.get_nerf_models(device="cpu")
    assert c is not f


def test_cpu_tensors_are_rejected(D):
    with pytest.raises(RuntimeError):
        D.raw2outputs(torch.rand(2, 8, 3), torch.rand(2, 8, 1), torch.rand(2, 8).sort(-1).values)
    m = D.OptimizedVanillaNeRF()
    with pytest.raises(RuntimeError):
        m(torch.zeros(4, 3), torch.zeros(4, 3))


def test_config_loader(D, tmp_path):
    for name in ("base_cfg", "blender_scene", "grass_scene", "real_scene"):
        p = os.path.join("/root/reference/cfg", name + ".yml")
        if os.path.exists(p):
            c = D.load_config(p)
            assert c.checkpoint.endswith("checkpoints/nerf_final.pth") and "${" not in c.checkpoint
            assert c.N_samples == 64 and c.N_importance == 128 and c.hidden_dim == 256
    y = tmp_path / "c.yml"
    y.write_text('scene_name: "s"\noutput_dir: "./o/${scene_name}"\ncheckpoint_dir: "${output_dir}/ck"\nn: 3\nm: ${n}\nsub:\n  a: [1, 2]\n')
    c = D.load_config(str(y))
    assert c.checkpoint_dir == "./o/s/ck" and c.m == 3 and c.sub.a == [1, 2]


def test_onecycle_tables_match_oracle(D):
    from oracle import nerf_oracle as O
    lrs, b1s = D.onecycle_tables(200, 5e-4)
    for s in (0, 1, 19, 20, 100, 199):
        lr, b1 = O.onecycle(s, 200, 5e-4)
        assert lrs[s] == pytest.approx(lr, rel=1e-12) and b1s[s] == pytest.approx(b1, rel=1e-12)


def test_bench_aux_deadline_prints_headline_once_and_exits_zero(tmp_path):
    """bench.py's auxiliary sections run under a deadline: if one never returns, rank 0 still prints ONE JSON line that
    carries the headline numbers and names the section, and the process exits 0; a run that finishes in time prints the
    line itself and the timer stays quiet."""
    import json, subprocess, sys, textwrap
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = textwrap.dedent("""
        import sys, time, json
        sys.path.insert(0, %r)
        import bench
        line = {"metric": bench.METRIC, "value": 1.0, "unit": bench.UNIT, "extract_vol": None}
        g = bench.AuxDeadline(float(sys.argv[1]), int(sys.argv[2]), line)
        g.enter("extract_vol")
        if sys.argv[3] == "hang":
            time.sleep(30)
            sys.exit(7)
        line["extract_vol"] = {"ms": 1.0}
        if g.finish() and int(sys.argv[2]) == 0:
            print(json.dumps(line), flush=True)
        time.sleep(1.0)
    """ % root)
    script = tmp_path / "deadline.py"
    script.write_text(code)
    hung = subprocess.run([sys.executable, str(script), "0.5", "0", "hang"], capture_output=True, text=True, timeout=120)
    assert hung.returncode == 0
    lines = [l for l in hung.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    got = json.loads(lines[0])
    assert got["value"] == 1.0 and got["aux_timeout"]["section"] == "extract_vol" and got["extract_vol"] is None
    quiet = subprocess.run([sys.executable, str(script), "0.5", "1", "hang"], capture_output=True, text=True, timeout=120)
    assert quiet.returncode == 0 and quiet.stdout.strip() == ""          # other ranks leave without printing
    done = subprocess.run([sys.executable, str(script), "0.5", "0", "ok"], capture_output=True, text=True, timeout=120)
    lines = [l for l in done.stdout.splitlines() if l.strip()]
    assert done.returncode == 0 and len(lines) == 1 and "aux_timeout" not in json.loads(lines[0])
