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
