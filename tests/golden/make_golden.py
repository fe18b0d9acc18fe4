#!/usr/bin/env python
"""Extract the known-answer vectors of the reference's own tests into tests/golden/reference_kats.json.

Run in the build container (needs /root/reference, which the GPU box does not have):
    python tests/golden/make_golden.py
Only numeric literals are taken (test DATA); no reference code is copied.  Each entry records the
reference file and test function it came from.  Numeric list literals of a test function are stored
in source order; tests/test_oracle_kat.py indexes them by position and cites the line.
"""
import json
import os
import re

REF = os.environ.get("BF_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_kats.json")

NUM = r"[-+]?(?:\d+\.?\d*(?:[eE][-+]?\d+)?|\.\d+(?:[eE][-+]?\d+)?)"


def functions(path):
    src = open(os.path.join(REF, path)).read()
    parts = re.split(r"^def (\w+)\(", src, flags=re.M)
    return {parts[i]: parts[i + 1] for i in range(1, len(parts) - 1, 2)}


def numeric_lists(body, min_len=3):
    """All innermost [...] literals holding only numbers, flattened per outermost bracket group."""
    out = []
    depth, start = 0, None
    for i, ch in enumerate(body):
        if ch == "[":
            if depth == 0:
                start = i
            depth += 1
        elif ch == "]":
            depth -= 1
            if depth == 0 and start is not None:
                text = body[start:i + 1]
                if re.fullmatch(r"[\[\]\s,eE\d.+\-]*", text):
                    vals = [float(x) for x in re.findall(NUM, text)]
                    if len(vals) >= min_len:
                        out.append(vals)
                start = None
    return out


def main():
    g = {}
    # TEA (src/libcore/tests/test_random.py:6-27): assert sample_tea_floatXX(a, b, 4) == value
    f = functions("src/libcore/tests/test_random.py")
    for name, key in (("test01_tea_float32", "tea_float32"), ("test02_tea_float64", "tea_float64")):
        g[key] = {"source": "src/libcore/tests/test_random.py:" + name,
                  "cases": [[int(a), int(b), int(r), float(v)] for a, b, r, v in
                            re.findall(r"sample_tea_float\d+\((\d+), (\d+), (\d+)\) == (" + NUM + ")", f[name])]}
    # microfacet tables (src/librender/tests/test_microfacet.py)
    f = functions("src/librender/tests/test_microfacet.py")
    for name in ("test02_eval_pdf_beckmann", "test03_smith_g1_beckmann", "test04_sample_beckmann", "test03_smith_g1_ggx", "test05_sample_ggx"):
        if name in f:
            g["microfacet/" + name] = {"source": "src/librender/tests/test_microfacet.py:" + name, "lists": numeric_lists(f[name], 6)}
    # fresnel conductor (src/librender/tests/test_fresnel.py:63-76)
    f = functions("src/librender/tests/test_fresnel.py")
    if "test03_fresnel_conductor" in f:
        g["fresnel_conductor_body"] = {"source": "src/librender/tests/test_fresnel.py:test03_fresnel_conductor",
                                       "lists": numeric_lists(f["test03_fresnel_conductor"], 2), "text_sha": len(f["test03_fresnel_conductor"])}
    # mesh fixtures used by test_mesh.py:257-298 and test_kdtrees.py:60-91, stored as arrays (data files)
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    from beifong_b200 import meshes
    p, fc, _ = meshes.load_obj(os.path.join(REF, "resources/data/common/meshes/rectangle.obj"))
    g["rectangle_obj"] = {"source": "resources/data/common/meshes/rectangle.obj", "positions": p.reshape(-1).tolist(), "faces": fc.reshape(-1).tolist()}
    p, fc, _ = meshes.load_ply(os.path.join(REF, "resources/data/common/meshes/bunny_lowres.ply"))
    g["bunny_lowres_ply"] = {"source": "resources/data/common/meshes/bunny_lowres.ply", "positions": [float(x) for x in p.reshape(-1)],
                             "faces": fc.reshape(-1).tolist()}
    json.dump(g, open(OUT, "w"), indent=0)
    print("wrote", OUT, {k: (len(v.get("cases", v.get("lists", v.get("faces", [])))) ) for k, v in g.items()})


if __name__ == "__main__":
    main()
