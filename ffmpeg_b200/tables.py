"""Python view of the generated constant tables in csrc/vvc_tables.h (host-side derivations only)."""
import os
import re

import numpy as np

_H = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc", "vvc_tables.h")
_CACHE = {}
_CT = {"int8_t": np.int8, "uint8_t": np.uint8, "int16_t": np.int16, "uint16_t": np.uint16}


def table(name):
    if name not in _CACHE:
        with open(_H) as f:
            src = f.read()
        m = re.search(r"VVC_TABLE_QUAL\s+(\w+)\s+%s\[(\d+)\]\s*=\s*\{([^}]*)\}" % re.escape(name), src)
        if not m:
            raise KeyError(name)
        a = np.array([int(v) for v in m.group(3).replace("\n", " ").split(",") if v.strip()], _CT[m.group(1)])
        assert a.size == int(m.group(2))
        _CACHE[name] = a
    return _CACHE[name]


def __getattr__(name):
    if name == "LFNST_TR_SET":
        return table("VVC_LFNST_TR_SET")
    raise AttributeError(name)
