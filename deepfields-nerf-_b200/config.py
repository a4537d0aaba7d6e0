"""Loader for the reference's cfg/*.yml files: flat YAML namespace with OmegaConf-style ${key} / ${a.b}
interpolation (cfg/base_cfg.yml:3-6,42-46).  A self-contained loader: omegaconf is not required (nor used)."""
from __future__ import annotations

import re
from types import SimpleNamespace

import yaml

_PAT = re.compile(r"\$\{([^}]+)\}")


class Config(SimpleNamespace):
    def __getitem__(self, k):
        return getattr(self, k)

    def get(self, k, default=None):
        return getattr(self, k, default)


def _lookup(root: dict, dotted: str):
    cur = root
    for part in dotted.split("."):
        cur = cur[part]
    return cur


def _resolve(value, root, depth=0):
    if depth > 32:
        raise ValueError("config interpolation too deep (cycle?)")
    if isinstance(value, str):
        m = _PAT.fullmatch(value)
        if m:  # whole-string reference keeps the referenced type
            return _resolve(_lookup(root, m.group(1)), root, depth + 1)
        return _PAT.sub(lambda mm: str(_resolve(_lookup(root, mm.group(1)), root, depth + 1)), value)
    if isinstance(value, dict):
        return {k: _resolve(v, root, depth) for k, v in value.items()}
    if isinstance(value, list):
        return [_resolve(v, root, depth) for v in value]
    return value


def _wrap(value):
    if isinstance(value, dict):
        return Config(**{k: _wrap(v) for k, v in value.items()})
    return value


def load_config(path: str) -> Config:
    with open(path) as fh:
        raw = yaml.safe_load(fh) or {}
    return _wrap(_resolve(raw, raw))
