"""TEST INFRASTRUCTURE ONLY (see oracle/pf_oracle.h): CPU restatement of the viewer's colour-map fragment shader,
reference gl.h:221-283, in numpy float32 (GLSL `float`), followed by the UNORM8 framebuffer conversion
(round to nearest). Parity unpinned by the reference (it ships no image fixtures, and the GL path is unavailable on
this box): the shader body restated statement by statement is the oracle.
"""
import numpy as np

PI = np.float32(3.14159265359)                                      # gl.h:232


def shade(values, min_val, max_val):
    """values: any float array -> (..., 4) uint8 RGBA."""
    f32 = np.float32
    val = np.asarray(values).astype(f32)                            # gl.h:396: the texture is uploaded as float
    lo, hi = f32(min_val), f32(max_val)
    r = np.zeros(val.shape, f32); g = np.zeros(val.shape, f32); b = np.zeros(val.shape, f32)
    with np.errstate(invalid="ignore", divide="ignore"):
        nan = np.isnan(val)                                         # gl.h:242-245
        below = ~nan & (val < lo)                                   # gl.h:246-252
        above = ~nan & ~below & (val > hi)                          # gl.h:253-258
        inside = ~nan & ~below & ~above
        d_below = (f32(1) - np.arctan(lo - val, dtype=f32) / PI * f32(2)) * f32(0.3)
        d_above = np.arctan(val - lo, dtype=f32) / PI * f32(2) * f32(0.3) + f32(0.7)
        v = np.minimum(np.maximum(val, lo), hi - f32(0.0001))       # gl.h:263
        d = hi - lo
        v = np.where(d == 0, f32(0.5), (v - lo) / d).astype(f32)
        m = f32(0.25)
        num = np.floor(v / m).astype(f32)
        s = ((v - num * m) / m).astype(f32)
    one, zero = np.ones(val.shape, f32), np.zeros(val.shape, f32)
    seg = np.where(np.isfinite(num), num, -1).astype(np.int64)
    table = {0: (zero, s, one), 1: (zero, one, one - s), 2: (s, one, zero), 3: (one, one - s, zero)}   # gl.h:271-276
    for k, (rr, gg, bb) in table.items():
        sel = inside & (seg == k)
        r = np.where(sel, rr, r); g = np.where(sel, gg, g); b = np.where(sel, bb, b)
    for sel, dd in ((below, d_below), (above, d_above)):
        r = np.where(sel, dd, r); g = np.where(sel, dd, g); b = np.where(sel, dd, b)
    r = np.where(nan, one, r); g = np.where(nan, zero, g); b = np.where(nan, one, b)

    def unorm8(c):
        c = np.clip(c.astype(f32), f32(0), f32(1))
        return np.floor(c * f32(255) + f32(0.5)).astype(np.uint8)

    out = np.empty(val.shape + (4,), np.uint8)
    out[..., 0], out[..., 1], out[..., 2], out[..., 3] = unorm8(r), unorm8(g), unorm8(b), 255
    return out
