"""TEST INFRASTRUCTURE ONLY (see oracle/oracle.cpp header): numpy restatement of HDRADC::bitmap(raw=False).

Follows src/adcs/hdradc.cpp:219-275 and the scalar StructConverter src/libcore/struct.cpp:1750-1830: the source
field with index 2 carries Struct::Flags::Weight and is left out of the target; per pixel inv_weight = 1.f / weight
(fp32), every target channel = source value * inv_weight, then stored in the component format (float16 is a
round-to-nearest-even conversion, float32 a copy).  Parity unpinned against the reference binary (Enoki is absent, the
Bitmap class does not build here); pinned only by the arithmetic above.
"""
import numpy as np


def develop(raw, component_format="float16"):
    raw = np.asarray(raw, dtype=np.float32)
    c = raw.shape[-1]
    keep = [i for i in range(c) if i != 2]
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        inv_weight = (np.float32(1.0) / raw[..., 2:3]).astype(np.float32)
        out = (raw[..., keep] * inv_weight).astype(np.float32)
        if component_format == "float16":
            return out.astype(np.float16)
    if component_format == "float32":
        return out
    raise ValueError(component_format)
