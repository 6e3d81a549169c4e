"""oracle/render_oracle.py -- TEST INFRASTRUCTURE, not product code.

NumPy restatement of the reference GUI's DrawParticles / CreateCamera (ref src/main.c:198-250) as a
software painter: particles in array order, each a filled disc centred on (int)pos (ref :244-245) of
radius max(radius, min_radius) (ref :246), colour by mass class (ref :227-235, colours :20-23), camera
transform screen = (world - target) * zoom + (W/2, H/2) (raylib Camera2D with offset at the window centre,
ref :199-202).  A pixel belongs to a disc when its CENTRE is inside.  The reference rasterises through
OpenGL (a 36-segment triangle fan), which cannot run here ("unavailable on this box"), so this restatement
-- not the reference -- defines the image; parity of the CUDA splat against it is bit-exact, parity against
the reference GUI is unpinned.  All arithmetic is FP32, one rounding per operation, like the kernel.
"""
import numpy as np

F = np.float32
BG = (22, 22, 22, 255)       # ref main.c:20
CC = (222, 222, 222, 255)    # ref main.c:21 galaxy core
NP_ = (175, 195, 175, 255)   # ref main.c:22 particle
EP = (145, 145, 233, 255)    # ref main.c:23 massless particle

# include/galaxy.h:21-25: MIN_GC_MASS = (4 * PI * GC_DENSITY / 3) * R * R * R with R = GC_MIN_R, in float
MIN_GC_MASS = F(F(F(F(F(F(4.0) * F(3.1415927)) * F(30.0)) / F(3.0)) * F(200.0)) * F(200.0)) * F(200.0)


def fit_camera(ps, width, height):
    """ref main.c:198-225 CreateCamera -> (target_x, target_y, zoom)."""
    minx, maxx = ps["px"].min(), ps["px"].max()
    miny, maxy = ps["py"].min(), ps["py"].max()
    zx = F(F(0.9) * F(width)) / F(maxx - minx)
    zy = F(F(0.9) * F(height)) / F(maxy - miny)
    zoom = min(zx, zy) if (zx < 1 or zy < 1) else F(1.0)
    return F(F(minx + maxx) * F(0.5)), F(F(miny + maxy) * F(0.5)), F(zoom)


def color_for_mass(mass):
    if mass <= 0:
        return EP
    if mass < MIN_GC_MASS:
        return NP_
    return CC


def render(ps, width, height, camera, min_radius=0.0):
    tx, ty, zoom = (F(v) for v in camera)
    img = np.empty((height, width, 4), dtype=np.uint8)
    img[:] = BG
    xs = (np.arange(width, dtype=np.float32) + F(0.5))
    ys = (np.arange(height, dtype=np.float32) + F(0.5))
    for p in ps:                                           # painter's order = array order
        wx, wy = F(np.trunc(p["px"])), F(np.trunc(p["py"]))
        sx = F(F(F(wx - tx) * zoom) + F(0.5) * F(width))
        sy = F(F(F(wy - ty) * zoom) + F(0.5) * F(height))
        r = F(max(F(p["radius"]), F(min_radius)) * zoom)
        if not (r >= 0) or not np.isfinite(sx) or not np.isfinite(sy):
            continue
        x0, x1 = int(max(np.floor(sx - r - 1), 0)), int(min(np.ceil(sx + r + 1), width - 1))
        y0, y1 = int(max(np.floor(sy - r - 1), 0)), int(min(np.ceil(sy + r + 1), height - 1))
        if x1 < x0 or y1 < y0:
            continue
        ddx = (xs[x0:x1 + 1] - sx).astype(np.float32)
        ddy = (ys[y0:y1 + 1] - sy).astype(np.float32)
        d2 = ((ddy * ddy).astype(np.float32)[:, None] + (ddx * ddx).astype(np.float32)[None, :]).astype(np.float32)
        # kernel order: ddx*ddx + ddy*ddy (addition is commutative: same rounding)
        inside = d2 <= F(r * r)
        img[y0:y1 + 1, x0:x1 + 1][inside] = color_for_mass(p["mass"])
    return img
