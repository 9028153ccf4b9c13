"""Minimal CfgNode: attribute access on a dict, merge_from_file (yaml), merge_from_list, clone, freeze/defrost.

dpvo/config.py:1 does `from yacs.config import CfgNode as CN`, builds defaults by attribute assignment and the drivers call
cfg.merge_from_file(<yaml>) and cfg.merge_from_list([...]) (dpvo_demo.py:171-175)."""
import copy

import yaml


class CfgNode(dict):
    def __init__(self, init=None):
        super().__init__()
        object.__setattr__(self, "_frozen", False)
        for k, v in (init or {}).items():
            self[k] = CfgNode(v) if isinstance(v, dict) and not isinstance(v, CfgNode) else v

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)

    def __setattr__(self, k, v):
        if object.__getattribute__(self, "_frozen"):
            raise AttributeError(f"attempted to set {k} on a frozen CfgNode")
        self[k] = v

    def _merge(self, other):
        for k, v in other.items():
            if k not in self:
                raise KeyError(f"non-existent config key: {k}")
            if isinstance(v, dict) and isinstance(self[k], CfgNode):
                self[k]._merge(v)
            else:
                old = self[k]
                if isinstance(old, float) and isinstance(v, int) and not isinstance(v, bool):
                    v = float(v)
                self[k] = v

    def merge_from_file(self, path):
        with open(path) as f:
            self._merge(yaml.safe_load(f) or {})

    def merge_from_other_cfg(self, other):
        self._merge(other)

    def merge_from_list(self, lst):
        assert len(lst) % 2 == 0
        for k, v in zip(lst[0::2], lst[1::2]):
            node = self
            *path, leaf = k.split(".")
            for p in path:
                node = node[p]
            if leaf not in node:
                raise KeyError(f"non-existent config key: {k}")
            if isinstance(v, str):
                try:
                    v = yaml.safe_load(v)
                except Exception:
                    pass
            old = node[leaf]
            if isinstance(old, float) and isinstance(v, int) and not isinstance(v, bool):
                v = float(v)
            node[leaf] = v

    def clone(self):
        return copy.deepcopy(self)

    def freeze(self):
        object.__setattr__(self, "_frozen", True)

    def defrost(self):
        object.__setattr__(self, "_frozen", False)

    def __deepcopy__(self, memo):
        out = CfgNode()
        for k, v in self.items():
            dict.__setitem__(out, k, copy.deepcopy(v, memo))
        return out
