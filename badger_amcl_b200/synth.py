"""Synthetic worlds, scans and particle clouds of the shapes SURVEY.md 8(d) names.

Plain numpy, deterministic (numpy Generator(PCG64) streams are platform independent). Used by tests/ and
bench.py to build inputs; nothing here is on the product's compute path and nothing here touches oracle/.
Cell convention is the reference's (occupancy_map.h:37-42): -1 free, 0 unknown, +1 occupied, stored as
cells[j, i] so that the flat index is i + j * size_x (occupancy_map.cpp:107-110).
"""
import math

import numpy as np

FREE, UNKNOWN, OCCUPIED = -1, 0, 1


def make_grid(size_x, size_y, seed=7, wall_every=200, clutter=0.001):
    """Rooms: border + axis-aligned walls every `wall_every` cells with door gaps + random clutter +
    a ring of unknown cells around a few rooms (SURVEY 8d 'common synthetic world')."""
    rng = np.random.default_rng(seed)
    cells = np.full((size_y, size_x), FREE, np.int8)
    jj, ii = np.meshgrid(np.arange(size_y), np.arange(size_x), indexing="ij")
    door = ((ii // 7 + jj // 11) % 5) == 0
    wall = ((ii % wall_every) == 0) | ((jj % wall_every) == 0)
    cells[wall & ~door] = OCCUPIED
    cells[0, :] = cells[-1, :] = OCCUPIED
    cells[:, 0] = cells[:, -1] = OCCUPIED
    cells[(rng.random((size_y, size_x)) < clutter) & (cells == FREE)] = OCCUPIED
    # unknown ring just inside a few rooms
    nrx, nry = max(1, size_x // wall_every), max(1, size_y // wall_every)
    for _ in range(max(1, (nrx * nry) // 8)):
        rx, ry = int(rng.integers(nrx)), int(rng.integers(nry))
        x0, y0 = rx * wall_every + 3, ry * wall_every + 3
        x1, y1 = min(size_x - 2, x0 + wall_every - 7), min(size_y - 2, y0 + wall_every - 7)
        if x1 <= x0 + 2 or y1 <= y0 + 2:
            continue
        for (ys, xs) in ((slice(y0, y0 + 1), slice(x0, x1)), (slice(y1, y1 + 1), slice(x0, x1)),
                         (slice(y0, y1), slice(x0, x0 + 1)), (slice(y0, y1), slice(x1, x1 + 1))):
            blk = cells[ys, xs]
            blk[blk == FREE] = UNKNOWN
    return cells


def world_to_map(x, y, size_x, size_y, resolution, origin=(0.0, 0.0)):
    """occupancy_map.cpp:90-98 (origin is float in the reference; pass float32-representable values)."""
    i = np.floor((np.asarray(x, np.float64) - np.float32(origin[0])) / resolution + 0.5).astype(np.int64) + size_x // 2
    j = np.floor((np.asarray(y, np.float64) - np.float32(origin[1])) / resolution + 0.5).astype(np.int64) + size_y // 2
    return i, j


def map_to_world(i, j, size_x, size_y, resolution, origin=(0.0, 0.0)):
    """occupancy_map.cpp:75-88"""
    return (np.float32(origin[0]) + (np.asarray(i) - size_x // 2) * resolution,
            np.float32(origin[1]) + (np.asarray(j) - size_y // 2) * resolution)


def raycast(cells, resolution, ox, oy, oa, max_range, origin=(0.0, 0.0)):
    """Vectorised restatement of the grid ray cast the scans are made with (same cell walk as
    occupancy_map.cpp:257-364: integer Bresenham, first non-free or off-map cell ends the ray)."""
    size_y, size_x = cells.shape
    oa = np.asarray(oa, np.float64)
    n = oa.shape[0]
    ox = np.broadcast_to(np.asarray(ox, np.float64), (n,))
    oy = np.broadcast_to(np.asarray(oy, np.float64), (n,))
    x0, y0 = world_to_map(ox, oy, size_x, size_y, resolution, origin)
    x1, y1 = world_to_map(ox + max_range * np.cos(oa), oy + max_range * np.sin(oa), size_x, size_y, resolution, origin)
    same = (x0 == x1) & (y0 == y1)
    steep = np.abs(y1 - y0) > np.abs(x1 - x0)
    a0, b0 = np.where(steep, y0, x0), np.where(steep, x0, y0)
    a1, b1 = np.where(steep, y1, x1), np.where(steep, x1, y1)
    da, db = np.abs(a1 - a0), np.abs(b1 - b0)
    astep, bstep = np.where(a0 < a1, 1, -1), np.where(b0 < b1, 1, -1)
    out = np.full(n, float(max_range))
    done = same.copy()
    nsteps = int(da.max()) + 1 if n else 0
    for s in range(0, nsteps + 1):
        live = ~done & (s <= da + 1)
        if not live.any():
            break
        a = a0 + s * astep
        b = b0 + bstep * ((2 * s * db + da) // np.maximum(2 * da, 1))
        cx, cy = np.where(steep, b, a), np.where(steep, a, b)
        inside = (cx >= 0) & (cx < size_x) & (cy >= 0) & (cy < size_y)
        state = np.where(inside, cells[np.clip(cy, 0, size_y - 1), np.clip(cx, 0, size_x - 1)], OCCUPIED)
        hit = live & (state != FREE)
        out[hit] = np.sqrt(((a - a0) ** 2 + (b - b0) ** 2)[hit].astype(np.float64)) * resolution
        done |= hit
    return out


def make_scan(cells, resolution, truth_pose, n_beams=1080, range_max=30.0, fov_deg=270.0, noise=0.02, dropouts=0.0,
              scanner_offset=(0.1, 0.0, 0.0), seed=11, origin=(0.0, 0.0)):
    """Ranges ray-cast from the ground-truth pose + N(0, noise); bearings evenly spread over the FOV.
    Readings at or beyond range_max are clamped to range_max (node_2d.cpp:534-560)."""
    rng = np.random.default_rng(seed)
    half = math.radians(fov_deg) / 2.0
    angles = np.linspace(-half, half, n_beams)
    th = truth_pose[2]
    sx = truth_pose[0] + scanner_offset[0] * math.cos(th) - scanner_offset[1] * math.sin(th)
    sy = truth_pose[1] + scanner_offset[0] * math.sin(th) + scanner_offset[1] * math.cos(th)
    ranges = raycast(cells, resolution, sx, sy, th + scanner_offset[2] + angles, range_max, origin)
    ranges = ranges + rng.normal(0.0, noise, n_beams)
    ranges = np.clip(ranges, 0.05, None)
    ranges[ranges >= range_max] = range_max
    if dropouts > 0:
        ranges[rng.random(n_beams) < dropouts] = range_max
    return ranges, angles


def free_cells(cells):
    """Free-space index list in the reference's i-major order (node_2d.cpp:324-335)."""
    jj, ii = np.nonzero(cells.T == FREE)  # transposed: iterate i outer, j inner
    return np.stack([jj, ii], axis=1)  # columns (i, j)


def uniform_free_poses(cells, resolution, n, seed=1, origin=(0.0, 0.0)):
    """Global-localisation cloud: uniform over free cells, theta uniform in [-pi, pi) (node.cpp:823-841)."""
    rng = np.random.default_rng(seed)
    fc = free_cells(cells)
    pick = fc[rng.integers(0, len(fc), n)]
    size_y, size_x = cells.shape
    x, y = map_to_world(pick[:, 0], pick[:, 1], size_x, size_y, resolution, origin)
    th = rng.random(n) * 2 * math.pi - math.pi
    return np.stack([x, y, th], axis=1).astype(np.float64)


def gaussian_poses(truth_pose, n, sigma_xy=0.5, sigma_th=0.26, seed=2):
    """Tracking cloud around the truth pose."""
    rng = np.random.default_rng(seed)
    p = np.empty((n, 3))
    p[:, 0] = truth_pose[0] + rng.normal(0, sigma_xy, n)
    p[:, 1] = truth_pose[1] + rng.normal(0, sigma_xy, n)
    p[:, 2] = truth_pose[2] + rng.normal(0, sigma_th, n)
    return p


def pick_truth_pose(cells, resolution, seed=5, origin=(0.0, 0.0)):
    """A free cell with some clearance, roughly mid-map."""
    rng = np.random.default_rng(seed)
    size_y, size_x = cells.shape
    for _ in range(10000):
        i = int(rng.integers(size_x // 4, 3 * size_x // 4))
        j = int(rng.integers(size_y // 4, 3 * size_y // 4))
        if (cells[max(0, j - 6):j + 7, max(0, i - 6):i + 7] == FREE).all():
            x, y = map_to_world(i, j, size_x, size_y, resolution, origin)
            return (float(x) + 0.013, float(y) - 0.007, float(rng.random() * 2 * math.pi - math.pi))
    raise RuntimeError("no free pose found")


class Drand48:
    """POSIX drand48 (the reference's only RNG, particle_filter.cpp:309): X <- (a X + c) mod 2^48."""
    A, C_, M = 0x5DEECE66D, 0xB, (1 << 48) - 1

    def __init__(self, seed=None):
        self.x = 0 if seed is None else (((seed & 0xFFFFFFFF) << 16) | 0x330E)

    def draw(self):
        self.x = (self.A * self.x + self.C_) & self.M
        return self.x / float(1 << 48)

    def state(self):
        return self.x


def reference_test_map():
    """The 100 x 150 map of the reference's own testOccupancyMapDistances (test/test_badger_amcl.cpp:131-154)."""
    cells = np.full((150, 100), FREE, np.int8)
    cells[3:12, 1] = UNKNOWN
    cells[10, 5:14] = OCCUPIED
    cells[15, 5:14] = OCCUPIED
    return cells


# ---------------------------------------------------------------------------------------------
# 3-D world: floor plane + wall slabs + clutter as occupied octree leaf keys
# ---------------------------------------------------------------------------------------------
def make_voxel_world(size_x=400, size_y=400, size_z=60, seed=3, wall_every=100, clutter=0.003):
    """Occupied leaf keys (n, 3) int32 of a room world: floor at k = 0, walls of full height every `wall_every`
    voxels with door gaps, random clutter in the lower third. Keys are centred on the origin in x / y."""
    rng = np.random.default_rng(seed)
    occ = np.zeros((size_z, size_y, size_x), bool)
    occ[0] = True
    jj, ii = np.meshgrid(np.arange(size_y), np.arange(size_x), indexing="ij")
    door = ((ii // 9 + jj // 13) % 4) == 0
    wall = (((ii % wall_every) == 0) | ((jj % wall_every) == 0)) & ~door
    wall[0, :] = wall[-1, :] = True
    wall[:, 0] = wall[:, -1] = True
    occ[:, wall] = True
    low = max(1, size_z // 3)
    occ[1:low] |= rng.random((low - 1, size_y, size_x)) < clutter
    k, j, i = np.nonzero(occ)
    keys = np.stack([i - size_x // 2, j - size_y // 2, k], axis=1).astype(np.int32)
    return keys


def make_cloud(keys, resolution, truth_pose, n_points=4096, mount=(0.2, 0.0, 1.5), max_range=20.0, seed=4, noise=0.01):
    """Sensor-frame cloud (n, 3) float32: occupied voxel centres seen from the sensor at the truth pose (within
    max_range, no occlusion test -- the scorer does not need one), sub-sampled to n_points, plus noise."""
    rng = np.random.default_rng(seed)
    c = (keys.astype(np.float64) + 0.5) * resolution
    x, y, th = truth_pose
    cs, sn = math.cos(th), math.sin(th)
    sx = x + cs * mount[0] - sn * mount[1]
    sy = y + sn * mount[0] + cs * mount[1]
    sz = mount[2]
    d = c - np.array([sx, sy, sz])
    near = np.nonzero((d ** 2).sum(1) < max_range ** 2)[0]
    pick = near[rng.permutation(len(near))[:n_points]]
    pick.sort()
    d = d[pick]
    # world -> sensor frame (yaw only; the mount's small pitch is ignored by the generator)
    px = cs * d[:, 0] + sn * d[:, 1]
    py = -sn * d[:, 0] + cs * d[:, 1]
    pts = np.stack([px, py, d[:, 2]], axis=1) + rng.normal(0.0, noise, (len(d), 3))
    if len(pts) < n_points:  # small worlds: repeat with fresh noise
        reps = -(-n_points // max(1, len(pts)))
        pts = np.concatenate([pts + rng.normal(0.0, noise, pts.shape) for _ in range(reps)])[:n_points]
    return np.ascontiguousarray(pts, np.float32)


def rpy_quaternion(roll, pitch, yaw):
    """tf2::Quaternion::setRPY (x, y, z, w)"""
    hr, hp, hy = roll * 0.5, pitch * 0.5, yaw * 0.5
    cr, sr, cp, sp, cy, sy = math.cos(hr), math.sin(hr), math.cos(hp), math.sin(hp), math.cos(hy), math.sin(hy)
    return (sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy,
            cr * cp * cy + sr * sp * sy)
