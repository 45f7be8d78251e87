#!/usr/bin/env python3
"""Extracts the pure virtual member functions of the interfaces the plugin derives from
(reference include/hpp/core/{obstacle-user,nearest-neighbor,config-validation,path-validation}.hh) into
tests/golden/plugin_virtuals.json.  Run in the development container (needs /root/reference)."""
import json
import os
import re
import sys

REF = "/root/reference/include/hpp/core"
WANT = {"ObstacleUserInterface": "obstacle-user.hh", "NearestNeighbor": "nearest-neighbor.hh",
        "ConfigValidation": "config-validation.hh", "PathValidation": "path-validation.hh"}


def pure_virtuals(path, cls):
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    m = re.search(r"class\s+(?:HPP_CORE_DLLAPI\s+)?%s\b[^;{]*\{" % cls, src)
    depth, i = 1, m.end()
    while depth and i < len(src):
        depth += {"{": 1, "}": -1}.get(src[i], 0)
        i += 1
    body = src[m.end():i]
    out = []
    for decl in re.findall(r"virtual\s+([^;{]*?)=\s*0\s*;", body, flags=re.S):
        decl = " ".join(decl.split())
        name = re.search(r"(\w+)\s*\(", decl).group(1)
        nargs = 0 if re.search(r"\(\s*\)", decl) else decl[decl.index("("):].count(",") + 1
        out.append({"name": name, "nargs": nargs, "decl": decl})
    return out


def extract():
    return {cls: pure_virtuals(os.path.join(REF, f), cls) for cls, f in WANT.items()}


if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "plugin_virtuals.json")
    json.dump(extract(), open(out, "w"), indent=1)
    print("wrote", out)
