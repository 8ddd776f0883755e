"""TEST INFRASTRUCTURE -- CPU oracle of the headless rasteriser (csrc/render.cu).  Only tests/, smoke() and
bench.py's CPU legs may import this.

numpy restatement of the fixed-function raster the reference performs through pyglet / OpenGL
(onpolicy/envs/mpe/environment.py:263-412, rendering.py:44-112,223-250): white clear; per entity (agents, then
landmarks) a filled `res`-gon of radius entity.size blended SRC_ALPHA / ONE_MINUS_SRC_ALPHA, then its outline with
colour and alpha halved (rendering.py:236-241); communication circles after all entities (environment.py:367-374).

Parity UNPINNED against the reference's pixels: its GL path cannot run on this box (no graphics stack), and GL leaves
smooth-line coverage to the implementation.  The rules fixed here (and in the kernel): pixel-centre sampling, outline
= centre within line_width/2 of an edge segment, blend results rounded to 1/255 like an RGBA8 colour buffer.
All arithmetic float64 in the kernel's operation order, so the CUDA image must match bit for bit.
"""
import math

import numpy as np


def unit_circle(res=30):
    """(cos, sin)(2 pi i / res) exactly as rendering.make_circle computes them (rendering.py:244-246)."""
    return np.array([[math.cos(2 * math.pi * i / res), math.sin(2 * math.pi * i / res)] for i in range(res)],
                    dtype=np.float64)


def primitives(pos, size, rgba, comm_dim=None, comm=None):
    """Draw list of one env: pos [E,2] f64, size [E], rgba [E,4] (float32 values), comm_dim [E] ints,
    comm [E] utterance id or -1.  Returns [(cx, cy, r, (r,g,b,a))]."""
    E = len(size)
    out = []
    for e in range(E):
        out.append((float(pos[e][0]), float(pos[e][1]), float(np.float32(size[e])),
                    tuple(float(np.float32(c)) for c in rgba[e])))
    for e in range(E):
        dc = 0 if comm_dim is None else int(comm_dim[e])
        if dc <= 0:
            continue
        es = float(np.float32(size[e]))
        cs = es / float(dc)
        said = -1 if comm is None else int(comm[e])
        for ci in range(dc):
            col = 0.0 if said == ci else 1.0
            out.append((out[e][0] + ((float(ci) * cs) * 2.0 - es + cs), out[e][1], cs, (col, col, col, 1.0)))
    return out


def _quant(v):
    return np.floor(v * 255.0 + 0.5) / 255.0


def rasterise(prims, width=700, height=700, cam_range=2.0, line_width=2.0, res=30):
    """uint8 [height, width, 3], row 0 = top."""
    W, H = int(width), int(height)
    cam = float(cam_range)
    sx, sy = W / (cam + cam), H / (cam + cam)
    hw = line_width * 0.5
    hw2 = hw * hw
    unit = unit_circle(res)
    col = np.ones((H, W, 3), dtype=np.float64)
    px = (np.arange(W, dtype=np.float64) + 0.5)[None, :]
    py = ((H - 1 - np.arange(H)).astype(np.float64) + 0.5)[:, None]
    for (cx, cy, r, c) in prims:
        pcx, pcy = (cx + cam) * sx, (cy + cam) * sy
        rmax = r * max(sx, sy) + hw
        # work on the bounding box only (the kernel's per-pixel circle test is a superset of every covered pixel)
        x_lo = max(int(math.floor(pcx - rmax - 1)), 0); x_hi = min(int(math.ceil(pcx + rmax + 1)), W)
        yb_lo = pcy - rmax - 1; yb_hi = pcy + rmax + 1                 # GL y
        r_lo = max(int(math.floor(H - 1 - yb_hi)), 0); r_hi = min(int(math.ceil(H - 1 - yb_lo)) + 1, H)
        if x_lo >= x_hi or r_lo >= r_hi:
            continue
        X = np.broadcast_to(px[:, x_lo:x_hi], (r_hi - r_lo, x_hi - x_lo))
        Y = np.broadcast_to(py[r_lo:r_hi, :], (r_hi - r_lo, x_hi - x_lo))
        dx, dy = X - pcx, Y - pcy
        near = (dx * dx + dy * dy) <= rmax * rmax
        inside = np.ones(X.shape, dtype=bool)
        edge = np.zeros(X.shape, dtype=bool)
        vx = ((cx + unit[:, 0] * r) + cam) * sx
        vy = ((cy + unit[:, 1] * r) + cam) * sy
        with np.errstate(invalid="ignore", divide="ignore"):
            for k in range(res):
                k1 = 0 if k + 1 == res else k + 1
                ex, ey = vx[k1] - vx[k], vy[k1] - vy[k]
                inv_len2 = 1.0 / (ex * ex + ey * ey)
                qx, qy = X - vx[k], Y - vy[k]
                cross = ex * qy - ey * qx
                inside &= cross >= 0.0
                t = (qx * ex + qy * ey) * inv_len2
                t = np.minimum(np.maximum(t, 0.0), 1.0)
                rx, ry = qx - t * ex, qy - t * ey
                edge |= (rx * rx + ry * ry) <= hw2
        inside &= near
        edge &= near
        sub = col[r_lo:r_hi, x_lo:x_hi]
        al, al2 = c[3], c[3] * 0.5
        for ch in range(3):
            v = sub[:, :, ch]
            v = np.where(inside, _quant(c[ch] * al + v * (1.0 - al)), v)
            v = np.where(edge, _quant((c[ch] * 0.5) * al2 + v * (1.0 - al2)), v)
            sub[:, :, ch] = v
    return np.floor(col * 255.0 + 0.5).astype(np.uint8)


def render_env(pos, size, rgba, comm_dim=None, comm=None, **kw):
    return rasterise(primitives(pos, size, rgba, comm_dim, comm), **kw)
