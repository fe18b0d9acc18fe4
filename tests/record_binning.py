"""numpy restatement of the tail of `receive_sample` for per-sample records (TEST INFRASTRUCTURE): where a sample lands
(`SignalBlock::put`, box filter: src/librender/signalblock.cpp:162-169 with `pos -= offset - border + 0.5`, `lo = ceil(pos - 0.5)`)
and which phase channel it goes to (`PhaseIntegrator::sample`, src/integrators/phase.cpp:117-124).  Written from the reference
text, independently of oracle/oracle.cpp and of the kernels; used to (a) check the device binning against a third
implementation and (b) isolate the samples whose bin differs between two implementations because an input differs by an ulp
("count them and show that excluding them restores 1e-3")."""
import numpy as np

TWO_PI = np.float32(6.28318530717958647692)


def pixel_of(rec, film_w, film_h, off_t=0, off_f=0):
    """Destination bin index (f * film_w + t) of each record, -1 outside the window."""
    px = rec["tf_t"].astype(np.float32) - np.float32(np.float32(off_t) + np.float32(0.5))
    py = rec["tf_f"].astype(np.float32) - np.float32(np.float32(off_f) + np.float32(0.5))
    with np.errstate(invalid="ignore"):
        cx, cy = np.ceil(px - np.float32(0.5)), np.ceil(py - np.float32(0.5))
        inside = (cx >= 0) & (cy >= 0) & (cx < film_w) & (cy < film_h)
    pix = np.full(len(rec), -1, np.int64)
    pix[inside] = cy[inside].astype(np.int64) * film_w + cx[inside].astype(np.int64)
    return pix


def phase_channel(rec, phase_bins):
    """AOV channel (0 .. phase_bins-1) the sample's value goes to, -1 if none (phase.cpp:117-124: the last matching bin wins)."""
    with np.errstate(invalid="ignore"):
        ph = np.fmod(rec["phase"].astype(np.float32), TWO_PI)
        ph = np.where(ph < 0, ph + TWO_PI, ph).astype(np.float32)
    bw = np.float32(TWO_PI / np.float32(phase_bins))
    ch = np.full(len(rec), -1, np.int64)
    for k in range(phase_bins):
        centre = np.float32(np.float64(bw) * (k + 0.5))
        with np.errstate(invalid="ignore"):
            hit = np.abs((ph - centre) / bw) < np.float32(0.5)
        ch[hit] = k
    return ch


def profile_from_records(rec, film_w, film_h, phase_bins, mask=None, off_t=0, off_f=0):
    """(film_h, film_w, 3 + phase_bins) float64 profile of the records selected by `mask`: Y = value, A = valid, W = 1."""
    pix = pixel_of(rec, film_w, film_h, off_t, off_f)
    ch = phase_channel(rec, phase_bins)
    sel = pix >= 0
    if mask is not None:
        sel &= mask
    C = 3 + phase_bins
    out = np.zeros((film_h * film_w, C), np.float64)
    val = np.where(rec["valid"] != 0, rec["value"].astype(np.float64), 0.0)
    np.add.at(out[:, 0], pix[sel], val[sel])
    np.add.at(out[:, 1], pix[sel], (rec["valid"][sel] != 0).astype(np.float64))
    np.add.at(out[:, 2], pix[sel], 1.0)
    return out.reshape(film_h, film_w, C), pix, ch


# ---- SignalBlock::put, filters wider than one bin -------------------------------------------------------------------------------
RAY_EPS = 1500.0 * 2.0 ** -24
FILTER_RESOLUTION = 31           # MTS_FILTER_RESOLUTION (include/mitsuba/core/rfilter.h)


def rfilter_eval(kind, x, radius):
    """The reference's reconstruction filters, from src/rfilters/{gaussian,tent,mitchell,catmullrom,lanczos}.cpp."""
    import math
    x = abs(x)
    if kind == 1:                                   # gaussian, stddev = radius / 4
        stddev = radius / 4.0
        alpha = -1.0 / (2.0 * stddev * stddev)
        return max(0.0, math.exp(alpha * x * x) - math.exp(alpha * radius * radius))
    if kind == 2:                                   # tent
        return max(0.0, 1.0 - x / radius)
    if kind in (3, 4):                              # mitchell (B = C = 1/3), catmullrom (B = 0, C = 1/2)
        B, C = (1.0 / 3.0, 1.0 / 3.0) if kind == 3 else (0.0, 0.5)
        x2, x3 = x * x, x * x * x
        if x < 1.0:
            r = (12 - 9 * B - 6 * C) * x3 + (-18 + 12 * B + 6 * C) * x2 + (6 - 2 * B)
        else:
            r = (-B - 6 * C) * x3 + (6 * B + 30 * C) * x2 + (-12 * B - 48 * C) * x + (8 * B + 24 * C)
        return r / 6.0 if x < 2.0 else 0.0
    if kind == 5:                                   # lanczos, lobes = radius
        if x < 2.0 ** -24:
            return 1.0
        if x > radius:
            return 0.0
        x1 = math.pi * x
        x2 = x1 / radius
        return (math.sin(x1) * math.sin(x2)) / (x1 * x2)
    raise ValueError(kind)


def wide_splat(size_x, size_y, kind, radius, px, py, off=(0, 0)):
    """One unit-valued sample splatted by SignalBlock::put's wide branch (src/librender/signalblock.cpp:107-160) over the discretised
    filter (rfilter.cpp:9-20, rfilter.h:62-65) into a block with its border, then cropped to the storage as ADC::put(block) does
    (signalblock.cpp:49-77).  Written from the reference text; compared with the oracle's splat in tests/test_oracle_kat.py."""
    import math
    table = [rfilter_eval(kind, radius * i / FILTER_RESOLUTION, radius) for i in range(FILTER_RESOLUTION)] + [0.0]
    scale = FILTER_RESOLUTION / radius
    border = int(math.ceil(radius - 0.5 - 2.0 * RAY_EPS))
    sx, sy = size_x + 2 * border, size_y + 2 * border
    pos = (px - (off[0] - border + 0.5), py - (off[1] - border + 0.5))
    lo = (max(math.ceil(pos[0] - radius), 0), max(math.ceil(pos[1] - radius), 0))
    hi = (min(math.floor(pos[0] + radius), sx - 1), min(math.floor(pos[1] + radius), sy - 1))
    n = int(math.ceil((radius - 2.0 * RAY_EPS) * 2.0))
    w = lambda x: table[min(int(abs(x * scale)), FILTER_RESOLUTION)]
    wx = [w(lo[0] - pos[0] + i) for i in range(n)]
    wy = [w(lo[1] - pos[1] + i) for i in range(n)]
    data = np.zeros((sy, sx))
    for yr in range(n):
        y = lo[1] + yr
        if y > hi[1]:
            continue
        for xr in range(n):
            x = lo[0] + xr
            if x > hi[0]:
                break
            data[y, x] += wy[yr] * wx[xr]
    return data[border:border + size_y, border:border + size_x]
