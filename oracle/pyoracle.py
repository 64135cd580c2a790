"""ctypes front-end of the CPU oracle (oracle/abm_oracle.cpp).

TEST INFRASTRUCTURE ONLY.  May be imported from tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs — never from the product
package (agents.jl_b200/).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None

METRICS = {"chebyshev": 0, "manhattan": 1, "euclidean": 2}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "abm_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        L = _lib
        L.orc_mod_f64.restype = C.c_double
        L.orc_mod_f64.argtypes = [C.c_double, C.c_double]
        L.orc_mod_f32.restype = C.c_float
        L.orc_mod_f32.argtypes = [C.c_float, C.c_float]
        L.orc_mod1.restype = C.c_int64
        L.orc_mod1.argtypes = [C.c_int64, C.c_int64]
        L.orc_grid_offsets.restype = C.c_int64
        L.orc_cs_create.restype = C.c_void_p
        L.orc_last_error.restype = C.c_char_p
        L.orc_last_error.argtypes = [C.c_void_p]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_grid_create.restype = C.c_void_p
        L.orc_cs_euclidean_distance.restype = C.c_double
        for name in ("orc_cs_nagents", "orc_cs_add_agents", "orc_cs_cell_ids", "orc_cs_grid_r",
                     "orc_cs_nearby_ids", "orc_cs_nearby_ids_agent", "orc_cs_pairs_all",
                     "orc_cs_nearest_neighbor", "orc_grid_add_agents", "orc_grid_nearby_positions",
                     "orc_grid_nearby_ids", "orc_grid_nearby_ids_agent"):
            getattr(L, name).restype = C.c_int64
    return _lib


def set_threads(n=0):
    """Threads of the synchronous flock step's gather phase (0 = all host cores).  The asynchronous,
    reference-faithful step is sequential like the reference's step!."""
    L = lib()
    L.orc_set_threads.restype = C.c_int
    return int(L.orc_set_threads(C.c_int(int(n))))


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def jl_mod(x, y, dtype=np.float64):
    if np.dtype(dtype) == np.float32:
        return float(lib().orc_mod_f32(C.c_float(x), C.c_float(y)))
    return float(lib().orc_mod_f64(x, y))


def grid_offsets(dims, periodic, metric, r, no_zero=False):
    D = len(dims)
    d = np.asarray(dims, dtype=np.int64)
    per = _periodic(periodic, D)
    n = lib().orc_grid_offsets(D, _p(d), _p(per), METRICS[metric], C.c_double(r), int(no_zero), None, C.c_int64(0))
    out = np.zeros((n, D), dtype=np.int64)
    lib().orc_grid_offsets(D, _p(d), _p(per), METRICS[metric], C.c_double(r), int(no_zero), _p(out), C.c_int64(n))
    return out


def _periodic(periodic, D):
    if isinstance(periodic, (bool, np.bool_)):
        periodic = (bool(periodic),) * D
    return np.asarray(periodic, dtype=np.uint8)


class OracleError(RuntimeError):
    pass


class ContinuousOracle:
    """Reference-semantics ContinuousSpace model (src/spaces/continuous.jl)."""

    def __init__(self, extent, spacing=None, periodic=True, dtype=np.float64):
        self.D = len(extent)
        self.dtype = np.dtype(dtype)
        ext = np.asarray(extent, dtype=np.float64)
        if spacing is None:
            spacing = float(ext.min()) / 20.0  # continuous.jl:95
        self.spacing = float(spacing)
        self.extent = ext
        self.periodic = _periodic(periodic, self.D)
        err = C.create_string_buffer(256)
        self.h = lib().orc_cs_create(self.D, _p(ext), C.c_double(self.spacing), _p(self.periodic),
                                     0 if self.dtype == np.float64 else 1, err, 256)
        if not self.h:
            raise OracleError(err.value.decode())
        self.h = C.c_void_p(self.h)
        dims = np.zeros(3, dtype=np.int64)
        lib().orc_cs_dims(self.h, _p(dims))
        self.dims = tuple(int(x) for x in dims[: self.D])

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orc_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def _err(self):
        return lib().orc_last_error(self.h).decode()

    def _arr(self, a):
        return np.ascontiguousarray(np.asarray(a, dtype=self.dtype))

    @property
    def n(self):
        return int(lib().orc_cs_nagents(self.h))

    def add_agents(self, pos, vel=None):
        pos = self._arr(pos).reshape(-1, self.D)
        vel = self._arr(vel).reshape(-1, self.D) if vel is not None else None
        rc = lib().orc_cs_add_agents(self.h, _p(pos), _p(vel), C.c_int64(pos.shape[0]))
        if rc < 0:
            raise OracleError(self._err())
        return int(rc)

    def add_agent(self, pos, vel=None):
        return self.add_agents([pos], [vel] if vel is not None else None)

    def state(self):
        n = self.n
        pos = np.zeros((n, self.D), dtype=self.dtype)
        vel = np.zeros((n, self.D), dtype=self.dtype)
        lib().orc_cs_get_state(self.h, _p(pos), _p(vel))
        return pos, vel

    def set_vel(self, vel):
        vel = self._arr(vel).reshape(-1, self.D)
        assert vel.shape[0] == self.n
        lib().orc_cs_set_vel(self.h, _p(vel))

    def pos2cell(self, pos):
        pos = self._arr(pos).reshape(-1, self.D)
        out = np.zeros((pos.shape[0], self.D), dtype=np.int32)
        lib().orc_cs_pos2cell(self.h, _p(pos), C.c_int64(pos.shape[0]), _p(out))
        return out

    def cell_ids(self, cell):
        c = np.asarray(cell, dtype=np.int64)
        n = lib().orc_cs_cell_ids(self.h, _p(c), None, C.c_int64(0))
        out = np.zeros(n, dtype=np.int64)
        lib().orc_cs_cell_ids(self.h, _p(c), _p(out), C.c_int64(n))
        return out

    def grid_r(self, pos, r):
        return int(lib().orc_cs_grid_r(self.h, _p(self._arr(pos)), C.c_double(r)))

    def nearby_ids(self, pos, r=1.0, search="approximate"):
        q = self._arr(pos)
        ex = int(search == "exact")
        n = lib().orc_cs_nearby_ids(self.h, _p(q), C.c_double(r), ex, None, C.c_int64(0))
        out = np.zeros(n, dtype=np.int64)
        lib().orc_cs_nearby_ids(self.h, _p(q), C.c_double(r), ex, _p(out), C.c_int64(n))
        return out

    def nearby_ids_agent(self, id, r=1.0, search="approximate"):
        ex = int(search == "exact")
        n = lib().orc_cs_nearby_ids_agent(self.h, C.c_int64(id), C.c_double(r), ex, None, C.c_int64(0))
        out = np.zeros(n, dtype=np.int64)
        lib().orc_cs_nearby_ids_agent(self.h, C.c_int64(id), C.c_double(r), ex, _p(out), C.c_int64(n))
        return out

    def nearby_counts_all(self, r, search="approximate"):
        out = np.zeros(self.n, dtype=np.int64)
        lib().orc_cs_nearby_counts_all(self.h, C.c_double(r), int(search == "exact"), _p(out))
        return out

    def interacting_pairs_all(self, r, search="exact"):
        ex = int(search == "exact")
        n = lib().orc_cs_pairs_all(self.h, C.c_double(r), ex, None, C.c_int64(0))
        out = np.zeros((n, 2), dtype=np.int64)
        lib().orc_cs_pairs_all(self.h, C.c_double(r), ex, _p(out), C.c_int64(n))
        return out

    def nearest_neighbor(self, id, r):
        j = int(lib().orc_cs_nearest_neighbor(self.h, C.c_int64(id), C.c_double(r)))
        return j or None

    def interacting_pairs_nearest(self, r, order=None):
        """true_pairs! (src/spaces/continuous.jl:454-497) restated literally: lists and linear scans, the
        scheduler's id order (default ByID).  Small populations only (quadratic like the reference)."""
        pos = self.state()[0]
        pairs, distances = [], []
        ids = range(1, int(self.n) + 1) if order is None else [int(i) for i in order]
        for a in ids:
            nn = self.nearest_neighbor(a, r)
            if nn is None:
                continue
            new_pair = (a, nn) if a < nn else (nn, a)
            if new_pair not in pairs:
                dist = self.euclidean_distance(pos[a - 1], pos[nn - 1])
                firsts = [p[0] for p in pairs]
                idx = firsts.index(new_pair[0]) if new_pair[0] in firsts else None
                if idx is None:
                    pairs.append(new_pair)
                    distances.append(dist)
                elif distances[idx] > dist:
                    pairs[idx] = new_pair
                    distances[idx] = dist
        to_remove = []
        flat = [i for p in pairs for i in p]
        for doubles in sorted({i for i in flat if flat.count(i) > 1}):
            firsts = [p[0] for p in pairs]
            lasts = [p[1] for p in pairs]
            if doubles in firsts:
                fidx, lidx = firsts.index(doubles), lasts.index(doubles)
                to_remove.append(lidx if distances[fidx] <= distances[lidx] else fidx)
            else:
                idxs = [i for i, v in enumerate(lasts) if v == doubles]
                keep = min(range(len(idxs)), key=lambda q: (distances[idxs[q]], q))
                del idxs[keep]
                to_remove.extend(idxs)
        dead = set(to_remove)
        return [p for i, p in enumerate(pairs) if i not in dead]


    def move_agent_dt(self, id, dt):
        if lib().orc_cs_move_agent_dt(self.h, C.c_int64(id), C.c_double(dt)):
            raise OracleError(self._err())

    def move_agent_to(self, id, pos):
        if lib().orc_cs_move_agent_to(self.h, C.c_int64(id), _p(self._arr(pos))):
            raise OracleError(self._err())

    def walk(self, id, direction):
        if lib().orc_cs_walk(self.h, C.c_int64(id), _p(self._arr(direction))):
            raise OracleError(self._err())

    def normalize_position(self, pos):
        p = self._arr(pos).copy()
        lib().orc_cs_normalize_position(self.h, _p(p))
        return p

    def euclidean_distance(self, a, b):
        return float(lib().orc_cs_euclidean_distance(self.h, _p(self._arr(a)), _p(self._arr(b))))

    def get_direction(self, a, b):
        out = np.zeros(self.D, dtype=self.dtype)
        lib().orc_cs_get_direction(self.h, _p(self._arr(a)), _p(self._arr(b)), _p(out))
        return out

    @staticmethod
    def _params(params):
        keys = ("speed", "cohere_factor", "separation", "separate_factor", "match_factor", "visual_distance")
        return np.asarray([float(params[k]) for k in keys], dtype=np.float64)

    def flock_step_sync(self, params, nsteps=1, search="approximate", per_agent=None, return_counts=False):
        p = self._params(params)
        pa = np.ascontiguousarray(per_agent, dtype=np.float64) if per_agent is not None else None
        cnt = np.zeros(self.n, dtype=np.int64) if return_counts else None
        rc = lib().orc_flock_step_sync(self.h, _p(p), _p(pa), int(search == "exact"), C.c_int64(nsteps), _p(cnt))
        if rc:
            raise OracleError(self._err())
        return cnt

    def flock_step_async(self, params, nsteps=1, search="approximate", per_agent=None, order=None):
        p = self._params(params)
        pa = np.ascontiguousarray(per_agent, dtype=np.float64) if per_agent is not None else None
        n = self.n
        o = np.ascontiguousarray(order, dtype=np.int64) if order is not None else None
        rc = lib().orc_flock_step_async(self.h, _p(p), _p(pa), int(search == "exact"), _p(o),
                                        C.c_int64(len(o) if o is not None else n), C.c_int64(nsteps))
        if rc:
            raise OracleError(self._err())

    def pairforce_step(self, radius, k, mass, dt, cutoff, nsteps=1):
        r = np.ascontiguousarray(radius, dtype=np.float64)
        kk = np.ascontiguousarray(k, dtype=np.float64)
        m = np.ascontiguousarray(mass, dtype=np.float64)
        rc = lib().orc_pairforce_step(self.h, _p(r), _p(kk), _p(m), C.c_double(dt), C.c_double(cutoff), C.c_int64(nsteps))
        if rc:
            raise OracleError(self._err())



def spatial_index(pos, extent, shape, dtype=np.float64):
    """get_spatial_index (src/spaces/continuous.jl:566-573): εs = spacesize[1:Dp] ./ size(property);
    idxs = floor.(Int, pos[1:Dp] ./ εs) .+ 1, in the space's element type.  Returns (n, Dp) 1-based indices."""
    pos = np.asarray(pos, dtype=dtype).reshape(-1, len(extent))
    Dp = len(shape)
    eps = np.asarray(extent[:Dp], dtype=dtype) / np.asarray(shape, dtype=dtype)
    return np.floor(pos[:, :Dp] / eps[None, :]).astype(np.int64) + 1


def elastic_collision(x1, v1, x2, v2, m1=1.0, m2=1.0):
    """elastic_collision!(a, b, f) (src/spaces/continuous.jl:333-366) on plain floats, same operation order.
    Returns (happened, v1_new, v2_new)."""
    import math

    r1 = (x1[0] - x2[0], x1[1] - x2[1])
    n = math.sqrt(r1[0] * r1[0] + r1[1] * r1[1]) ** 2
    if n == 0:
        return False, tuple(v1), tuple(v2)
    dv = (v1[0] - v2[0], v1[1] - v2[1])
    r2 = (x2[0] - x1[0], x2[1] - x1[1])
    dot = lambda a, b: a[0] * b[0] + a[1] * b[1]
    if m1 == math.inf and m2 == math.inf:
        return False, tuple(v1), tuple(v2)
    if m1 == math.inf:
        assert tuple(v1) == (0.0, 0.0), "An agent with ∞ mass cannot have nonzero velocity"
        if dot(r1, v2) <= 0:
            return False, tuple(v1), tuple(v2)
        v1 = (0.0, 0.0)
        f1, f2 = 0.0, 2.0
    elif m2 == math.inf:
        assert tuple(v2) == (0.0, 0.0), "An agent with ∞ mass cannot have nonzero velocity"
        if dot(r2, v1) <= 0:
            return False, tuple(v1), tuple(v2)
        v2 = (0.0, 0.0)
        f1, f2 = 2.0, 0.0
    else:
        if dot(dv, r2) <= 0:
            return False, tuple(v1), tuple(v2)
        f1 = 2 * m2 / (m1 + m2)
        f2 = 2 * m1 / (m1 + m2)
    q1, q2 = dot(dv, r1) / n, dot(dv, r2) / n
    return (True, tuple(v1[d] - f1 * q1 * r1[d] for d in range(2)), tuple(v2[d] + f2 * q2 * r2[d] for d in range(2)))


class GridOracle:
    """Reference-semantics GridSpace / GridSpaceSingle (src/spaces/grid_multi.jl, grid_single.jl)."""

    def __init__(self, dims, periodic=True, metric="chebyshev", single=False):
        self.D = len(dims)
        self.dims = tuple(int(x) for x in dims)
        self.periodic = _periodic(periodic, self.D)
        self.metric = metric
        self.single = bool(single)
        d = np.asarray(dims, dtype=np.int64)
        self.h = C.c_void_p(lib().orc_grid_create(self.D, _p(d), _p(self.periodic), METRICS[metric], int(single)))
        self.n = 0

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orc_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def _err(self):
        return lib().orc_last_error(self.h).decode()

    def add_agents(self, pos):
        p = np.ascontiguousarray(pos, dtype=np.int64).reshape(-1, self.D)
        rc = lib().orc_grid_add_agents(self.h, _p(p), C.c_int64(p.shape[0]))
        if rc < 0:
            raise OracleError(self._err())
        self.n = int(rc)
        return self.n

    def add_agent(self, pos):
        return self.add_agents([pos])

    def positions(self):
        out = np.zeros((self.n, self.D), dtype=np.int64)
        lib().orc_grid_get_pos(self.h, _p(out))
        return out

    def stored_single(self):
        out = np.zeros(int(np.prod(self.dims)), dtype=np.int64)
        lib().orc_grid_get_stored_single(self.h, _p(out))
        return out.reshape(self.dims[::-1]).T if self.D > 1 else out  # column-major → [x, y, ...]

    def move_agent_to(self, id, pos):
        if lib().orc_grid_move_agent_to(self.h, C.c_int64(id), _p(np.asarray(pos, dtype=np.int64))):
            raise OracleError(self._err())

    def walk(self, id, direction, ifempty=True):
        if lib().orc_grid_walk(self.h, C.c_int64(id), _p(np.asarray(direction, dtype=np.int64)), int(ifempty)):
            raise OracleError(self._err())

    def _collect(self, fn, first, r, width):
        n = fn(self.h, first, C.c_double(r), None, C.c_int64(0))
        out = np.zeros((n, width) if width > 1 else n, dtype=np.int64)
        fn(self.h, first, C.c_double(r), _p(out), C.c_int64(n))
        return out

    def nearby_positions(self, pos, r=1):
        o = np.asarray(pos, dtype=np.int64)
        out = self._collect(lib().orc_grid_nearby_positions, _p(o), r, self.D)
        return out.reshape(-1, self.D)

    def nearby_ids(self, pos, r=1):
        o = np.asarray(pos, dtype=np.int64)
        return self._collect(lib().orc_grid_nearby_ids, _p(o), r, 1)

    def nearby_ids_agent(self, id, r=1):
        return self._collect(lib().orc_grid_nearby_ids_agent, C.c_int64(id), r, 1)

    def schelling_happiness(self, group, min_to_be_happy):
        g = np.ascontiguousarray(group, dtype=np.int32)
        cnt = np.zeros(self.n, dtype=np.int32)
        mood = np.zeros(self.n, dtype=np.uint8)
        if lib().orc_schelling_happiness(self.h, _p(g), int(min_to_be_happy), _p(cnt), _p(mood)):
            raise OracleError(self._err())
        return cnt, mood

    def schelling_step_async(self, group, min_to_be_happy, seed=42, nsteps=1):
        g = np.ascontiguousarray(group, dtype=np.int32)
        mood = np.zeros(self.n, dtype=np.uint8)
        if lib().orc_schelling_step_async(self.h, _p(g), int(min_to_be_happy), _p(mood), C.c_uint64(seed), C.c_int64(nsteps)):
            raise OracleError(self._err())
        return mood


# ---------------------------------------------------------------------------
# randomwalk! — src/spaces/walk.jl:154-214 (grids), :389-408 (uniform_randomwalk!).  The reference draws from the
# model RNG; the counter stream below is the one the device rule is defined on (include/agents_b200.h,
# ab_randomwalk), so the oracle and the device see the same random numbers.
# ---------------------------------------------------------------------------
_M64 = (1 << 64) - 1


def _splitmix64_int(x):
    x = (x + 0x9E3779B97F4A7C15) & _M64
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
    return z ^ (z >> 31)


def rw_u01(seed, id, c):
    return (_splitmix64_int((seed ^ (((id << 16) + c) & _M64)) & _M64) >> 11) * 2.0 ** -53


def grid_offsets_at_radius(dims, periodic, metric, r):
    """offsets_at_radius(space, r) — grid_general.jl:119-136."""
    r0 = int(np.floor(r + 1)) - 1
    offs = grid_offsets(dims, periodic, metric, r)
    if metric == "manhattan":
        keep = np.abs(offs).sum(axis=1) == r0
    elif metric == "chebyshev":
        keep = np.abs(offs).max(axis=1) == r0
    else:
        raise OracleError("Random walks on a `GridSpace` with Euclidean metric are not defined.")
    return offs[keep]


def grid_randomwalk(o, ids, r, seed, ifempty=True, force_motion=False, literal=False):
    """randomwalk!(agent, model, r; ifempty, force_motion) agent after agent on a GridOracle (walk.jl:154-214).
    With force_motion the agent moves TO the chosen empty position (the documented behaviour); `literal=True`
    follows the reference's code as written instead: the chosen position is passed on to walk! as a direction
    (walk.jl:170-172 / :193-195), target = normalize_position(pos .+ choice), taken only if it is empty."""
    offs = grid_offsets_at_radius(o.dims, o.periodic, o.metric, r)
    L = len(offs)
    dims = np.asarray(o.dims, dtype=np.int64)
    per = _periodic(o.periodic, o.D)

    def norm(p):
        return tuple(int((x - 1) % s + 1) if pd else int(min(max(x, 1), s)) for x, s, pd in zip(p, dims, per))

    force = force_motion and (o.single or ifempty)
    occ = {}
    cur = [tuple(int(x) for x in p) for p in o.positions()]
    for p in cur:
        occ[p] = occ.get(p, 0) + 1
    for id in ids:
        p = cur[id - 1]
        if not force:
            j = min(int(np.floor(rw_u01(seed, id, 0) * L)), L - 1)
            tgt = norm(np.add(p, offs[j]))
            if (o.single or ifempty) and occ.get(tgt, 0) != 0:
                continue
        else:
            tgt = None
            for a in range(2 * L):
                j = min(int(np.floor(rw_u01(seed, id, a) * L)), L - 1)
                t = norm(np.add(p, offs[j]))
                if occ.get(t, 0) == 0:
                    tgt = t
                    break
            if tgt is None:
                empties = [t for t in (norm(np.add(p, b)) for b in offs) if occ.get(t, 0) == 0]
                if not empties:
                    continue
                tgt = empties[min(int(np.floor(rw_u01(seed, id, 2 * L) * len(empties))), len(empties) - 1)]
            if literal:
                tgt = norm(np.add(p, tgt))
                if occ.get(tgt, 0) != 0:
                    continue
        o.move_agent_to(id, tgt)
        occ[p] -= 1
        occ[tgt] = occ.get(tgt, 0) + 1
        cur[id - 1] = tgt


def continuous_randomwalk_directions(vel, ids, r, seed, dtype=np.float64):
    """direction of uniform_randomwalk!(agent, model, r) (walk.jl:389-408) for each id; r=None: the agent's speed."""
    T = np.dtype(dtype).type
    D = vel.shape[1]
    out = np.zeros((len(ids), D), dtype=dtype)
    for k, id in enumerate(ids):
        rr = float(np.sqrt(np.sum(vel[id - 1].astype(dtype) ** 2, dtype=dtype))) if r is None else float(r)
        if not rr > 0:
            raise OracleError("The displacement must be larger than 0.")
        v = np.zeros(D, dtype=dtype)
        for pr in range((D + 1) // 2):
            u1, u2 = rw_u01(seed, id, 2 * pr), rw_u01(seed, id, 2 * pr + 1)
            R, a = np.sqrt(-2.0 * np.log(1.0 - u1)), 6.283185307179586 * u2
            v[2 * pr] = T(R * np.cos(a))
            if 2 * pr + 1 < D:
                v[2 * pr + 1] = T(R * np.sin(a))
        nv = T(np.sqrt(np.sum(v * v, dtype=dtype)))
        if nv != 0:
            out[k] = ((v / nv).astype(np.float64) * rr).astype(dtype)
        else:
            out[k] = [(-1.0 if rw_u01(seed, id, 8 + d) < 0.5 else 1.0) * rr / np.sqrt(float(D)) for d in range(D)]
    return out


def rotate2(w, theta):
    """rotate(w::SVector{2}, θ) = Angle2d(θ) * w (walk.jl:332-337)."""
    c, s = np.cos(theta), np.sin(theta)
    return np.array([c * w[0] - s * w[1], s * w[0] + c * w[1]])


def _angle_axis(axis, ang, v):
    k = np.asarray(axis, dtype=np.float64)
    k = k / np.sqrt(np.sum(k * k))
    return v * np.cos(ang) + np.cross(k, v) * np.sin(ang) + k * np.dot(k, v) * (1.0 - np.cos(ang))


def rotate3(w, theta, phi):
    """rotate(w::SVector{3}, θ, ϕ) (walk.jl:347-357): θ about a vector normal to w, then ϕ about w."""
    w = np.asarray(w, dtype=np.float64)
    m = int(np.flatnonzero(w != 0)[0])
    n = (m + 1) % 3
    u = np.zeros(3)
    u[n] = w[m]
    u[m] = -w[n]
    a = _angle_axis(u, theta, w)
    return _angle_axis(w, phi, a)


def continuous_randomwalk_angle_directions(vel, ids, polar, azimuthal=None, r=None):
    """direction of randomwalk!(agent, model[, r]; polar, azimuthal) (walk.jl:236-330) for the given angles."""
    vel = np.asarray(vel, dtype=np.float64)
    out = np.zeros((len(ids), vel.shape[1]))
    for k, id in enumerate(ids):
        w = vel[id - 1]
        d = rotate2(w, polar[k]) if vel.shape[1] == 2 else rotate3(w, polar[k], azimuthal[k])
        if r is not None:
            d = d * (r / np.sqrt(np.sum(w * w)))
        out[k] = d
    return out
