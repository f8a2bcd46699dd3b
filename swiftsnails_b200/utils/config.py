"""`key: value` config files — Python twin of ``csrc/core/utils/config.h``.

ref: /root/reference/src/utils/ConfigParser.h:25-129. Same grammar: ``# comment``,
``import <path>`` (recursive), the FIRST ``:`` splits (so ``tcp://…`` values work), duplicate
keys keep the first value, a missing key raises (the reference aborts).  The complete key list
the framework reads is in SURVEY.md §2.12(5).
"""
from __future__ import annotations

import os


class ConfigParser:
    def __init__(self, path=None):
        self.items = {}
        if path:
            self.load_conf(path)

    def load_conf(self, path, _depth=0):
        if _depth > 16:
            raise RuntimeError("config import nesting too deep at %s" % path)
        base = os.path.dirname(os.path.abspath(path))
        with open(path) as f:
            for raw in f:
                self.parse_line(raw, base, _depth)
        return self

    def parse_line(self, raw, base=".", _depth=0):
        line = raw.split("#", 1)[0].strip()
        if not line:
            return False
        if line.startswith("import ") or line.startswith("import\t"):
            inc = line[6:].strip()
            cand = inc if os.path.isabs(inc) else os.path.join(base, inc)
            self.load_conf(cand if os.path.exists(cand) else inc, _depth + 1)
            return True
        key, _, val = line.partition(":")
        key = key.strip()
        if not key:
            raise ValueError("bad config line: %r" % raw)
        self.items.setdefault(key, val.strip())
        return True

    def has(self, key):
        return key in self.items

    def get(self, key, default=None):
        if key not in self.items:
            if default is None:
                raise KeyError("config key %r is missing" % key)
            return default
        return self.items[key]

    def to_int32(self, key, default=None):
        return int(self.get(key, None if default is None else str(default)))

    def to_float(self, key, default=None):
        return float(self.get(key, None if default is None else str(default)))

    def to_bool(self, key, default=None):
        return self.get(key, None if default is None else str(default)) in ("true", "True", "1")

    def to_string(self, key, default=None):
        return self.get(key, default)


_global = ConfigParser()


def global_config():
    return _global
