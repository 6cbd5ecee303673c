"""NumPy restatement of the frame renderer's arithmetic (float32, same formulas as
csrc/nbody_kernels.cuh splat_kernel / shade_kernel), used only by tests/test_gpu_render.py.
Camera and markers: src/renderers/GlutRenderer.cpp:19-30 (light), :98-141 (draw), :148 (projection)."""
import numpy as np

f32 = np.float32


def render_reference(pos, width, height, distance=800.0, angle_x=0.0, angle_y=0.0, radius=2.0):
    deg = f32(0.017453292519943295)
    cx, sx = f32(np.cos(f32(angle_x) * deg)), f32(np.sin(f32(angle_x) * deg))
    cy, sy = f32(np.cos(f32(angle_y) * deg)), f32(np.sin(f32(angle_y) * deg))
    fy = f32(0.5 * height / np.tan(0.5 * 45.0 * float(deg)))
    depth_buf = np.full((height, width), np.inf, np.float64)
    owner = np.full((height, width), -1, np.int64)
    img = np.zeros((height, width, 3), np.uint8)
    geom = []
    for k, (x, y, z) in enumerate(np.asarray(pos, f32)):
        x1 = x * cy + z * sy
        z1 = -x * sy + z * cy
        ex, ey, ez = x1, y * cx - z1 * sx, y * sx + z1 * cx - f32(distance)
        d = -ez
        if not (d > 1.0 and d < 5000.0):
            geom.append(None)
            continue
        px = f32(width * 0.5) + fy * ex / d
        py = f32(height * 0.5) - fy * ey / d
        rp = min(f32(fy * f32(radius) / d), f32(256.0))
        inv = f32(1.0) / max(rp, f32(0.5))
        geom.append((px, py, d, inv))
        x0, x1p = max(0, int(np.floor(px - rp))), min(width - 1, int(np.ceil(px + rp)))
        y0, y1p = max(0, int(np.floor(py - rp))), min(height - 1, int(np.ceil(py + rp)))
        for yy in range(y0, y1p + 1):
            for xx in range(x0, x1p + 1):
                u = (f32(xx + 0.5) - px) * inv
                v = (py - f32(yy + 0.5)) * inv
                q = f32(1.0) - u * u - v * v
                centre = xx == int(np.floor(px)) and yy == int(np.floor(py))
                if q < 0 and not centre:
                    continue
                zs = max(f32(d - f32(radius) * np.sqrt(max(q, f32(0.0)))), f32(0.0))
                if (zs, k) < (depth_buf[yy, xx], owner[yy, xx] if owner[yy, xx] >= 0 else 1 << 62):
                    depth_buf[yy, xx] = zs
                    owner[yy, xx] = k
    for yy in range(height):
        for xx in range(width):
            k = owner[yy, xx]
            if k < 0:
                continue
            px, py, d, inv = geom[k]
            u = (f32(xx + 0.5) - px) * inv
            v = (py - f32(yy + 0.5)) * inv
            w = np.sqrt(max(f32(0.0), f32(1.0) - u * u - v * v))
            ndotl = max(f32(0.0), (u + v + w) * f32(0.57735026919))
            lit = min(f32(1.0), f32(0.5) + ndotl)
            img[yy, xx] = (int(f32(255.0) * lit + f32(0.5)), int(f32(255.0) * min(f32(1.0), f32(0.5) * lit) + f32(0.5)), 0)
    return img, owner
