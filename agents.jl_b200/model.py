"""Host mirror of the Agents.jl surface that sits on the hot path.

Names and argument meaning follow the reference (Agents.jl v7.0.3):
  ContinuousSpace(extent; spacing, periodic)     src/spaces/continuous.jl:93-105
  GridSpace / GridSpaceSingle(dims; periodic, metric)   src/spaces/grid_multi.jl:68-86, grid_single.jl:33-45
  StandardABM(space; model_step!, properties)    src/core/model_standard.jl:134-177
  nearby_ids / interacting_pairs / nearest_neighbor / move_agent! / walk! / step!
Agent ids are 1-based and dense (the Vector-container rule,
src/core/model_accessing_API.jl:29-32).  All work happens in the CUDA library.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as L


class AgentsError(RuntimeError):
    """Raised where the reference would throw (ErrorException, BoundsError, ...)."""

    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def _periodic(periodic, D):
    if isinstance(periodic, (bool, np.bool_)):
        periodic = (bool(periodic),) * D
    if len(periodic) != D:
        raise ValueError("periodic must be a Bool or a D-tuple")
    return tuple(bool(p) for p in periodic)


class ContinuousSpace:
    """ContinuousSpace(extent; spacing = minimum(extent)/20, periodic = true) — continuous.jl:93-105."""

    kind = L.SPACE_CONTINUOUS

    def __init__(self, extent, spacing=None, periodic=True, dtype=np.float64):
        self.user_D = len(extent)
        self.extent = tuple(float(x) for x in extent)
        self.spacing = float(min(self.extent) / 20.0 if spacing is None else spacing)
        self.periodic = _periodic(periodic, self.user_D)
        # Float16 agents (test/continuous_space_tests.jl:606-637): the device computes and stores them in Float32 (every
        # Float16 is a Float32); columns are rounded to Float16 when they are read back.  The reference rounds after every
        # operation instead: identical for one operation per read-back (a move_agent!, a walk!), last-digit differences
        # for longer chains.
        self.user_dtype = np.dtype(dtype)
        self.dtype = np.dtype(np.float32) if self.user_dtype == np.float16 else self.user_dtype
        if self.dtype not in (np.dtype(np.float64), np.dtype(np.float32)):
            raise TypeError("eltype must be Float64, Float32 or Float16")
        self.metric = "euclidean"
        self.dims = None
        # A 1-D space is embedded in a 2-D one: a single non-periodic row of cells of height `spacing`, every agent at
        # mid-height with zero velocity in the second coordinate.  Distances, cells (first coordinate), grid_r and the
        # cells of the offset tables are those of the 1-D space (the second coordinate contributes exactly 0).
        self.D = max(self.user_D, 2)
        if self.user_D == 1:
            self.extent = (self.extent[0], self.spacing)
            self.periodic = (self.periodic[0], False)

    def _desc(self):
        d = L.SpaceDesc()
        d.kind, d.D = self.kind, self.D
        for i in range(self.D):
            d.extent[i] = self.extent[i]
            d.periodic[i] = int(self.periodic[i])
        d.spacing = self.spacing
        d.metric = L.METRICS["euclidean"]
        d.dtype = L.F64 if self.dtype == np.float64 else L.F32
        return d


class GridSpace:
    """GridSpace(d; periodic = true, metric = :chebyshev) — grid_multi.jl:68-86."""

    kind = L.SPACE_GRID_MULTI

    def __init__(self, dims, periodic=True, metric="chebyshev"):
        self.D = len(dims)
        self.dims = tuple(int(x) for x in dims)
        self.periodic = _periodic(periodic, self.D)
        if metric not in L.METRICS:
            raise ValueError("Unknown metric type")
        self.metric = metric
        self.dtype = np.dtype(np.int32)

    def _desc(self):
        d = L.SpaceDesc()
        d.kind, d.D = self.kind, self.D
        for i in range(self.D):
            d.periodic[i] = int(self.periodic[i])
            d.grid_dims[i] = self.dims[i]
        d.metric = L.METRICS[self.metric]
        d.dtype = L.I32
        return d


class GridSpaceSingle(GridSpace):
    """GridSpaceSingle(d; periodic, metric): at most one agent per position — grid_single.jl:33-45."""

    kind = L.SPACE_GRID_SINGLE


_NP2AB = {np.dtype(np.float64): L.F64, np.dtype(np.float32): L.F32, np.dtype(np.int32): L.I32,
          np.dtype(np.int64): L.I64, np.dtype(np.uint8): L.U8, np.dtype(np.bool_): L.U8}


class FlockRule:
    """`agent_step!` of examples/flock.jl:79-108 as a fused synchronous model step.
    Scalars are uniform; pass `per_agent=` (n×6 array: speed, cohere_factor, separation,
    separate_factor, match_factor, visual_distance) for per-bird fields (flock.jl:21-28)."""

    def __init__(self, speed=1.0, cohere_factor=0.1, separation=2.0, separate_factor=0.25, match_factor=0.04,
                 visual_distance=5.0, search="approximate", per_agent=None, kernel="auto"):
        self.p = dict(speed=speed, cohere_factor=cohere_factor, separation=separation,
                      separate_factor=separate_factor, match_factor=match_factor, visual_distance=visual_distance)
        self.search = search
        self.per_agent = per_agent
        self.kernel = {"auto": 0, "generic": 1, "tiled": 2, "fine": 2}[kernel]
        self._field = -1

    def _bind(self, model):
        if self.per_agent is not None and self._field < 0:
            self._field = model.add_column("_flock_params", np.asarray(self.per_agent, dtype=model.space.dtype))

    def __call__(self, model, nsteps=1, ids=None):
        self._bind(model)
        fp = L.FlockParams(search=L.SEARCH[self.search], per_agent_field=self._field, kernel=self.kernel, **self.p)
        if ids is None:
            model._ck(L.lib().ab_step_flock(model.h, C.byref(fp), int(nsteps)))
        else:       # schedule(model) returned a subset (Schedulers.Partially, a custom scheduler): schedulers.jl:68-120
            a = np.ascontiguousarray(ids, dtype=np.int64)
            model._ck(L.lib().ab_step_flock_active(model.h, C.byref(fp), int(nsteps), model._ptr(a), a.shape[0]))


class PairForceRule:
    """Force law + integrator of examples/celllistmap.jl:143-155, :176-191 (columns r, k, mass)."""

    def __init__(self, dt=0.001, cutoff=20.0):
        self.dt, self.cutoff = dt, cutoff

    def __call__(self, model, nsteps=1, ids=None):
        pp = L.PairForceParams(dt=self.dt, cutoff=self.cutoff, field_r=model.field_id("r"),
                               field_k=model.field_id("k"), field_mass=model.field_id("mass"))
        if ids is None:
            model._ck(L.lib().ab_step_pairforce(model.h, C.byref(pp), int(nsteps)))
        else:       # the forces of all particles (model_step!), the integrator (agent_step!) for the scheduled ones
            a = np.ascontiguousarray(ids, dtype=np.int64)
            model._ck(L.lib().ab_step_pairforce_active(model.h, C.byref(pp), int(nsteps), model._ptr(a), a.shape[0]))


class SchellingRule:
    """schelling_step! of docs/src/tutorial.jl:351-376, synchronous (columns group::Int32, mood::UInt8)."""

    def __init__(self, min_to_be_happy=3, relocate=True, seed=42):
        self.min_to_be_happy, self.relocate, self.seed = min_to_be_happy, relocate, seed
        self.last_moved = 0

    def __call__(self, model, nsteps=1, ids=None):
        sp = L.SchellingParams(min_to_be_happy=self.min_to_be_happy, relocate=int(self.relocate), seed=self.seed)
        moved = C.c_int64(0)
        if ids is None:
            model._ck(L.lib().ab_step_schelling(model.h, C.byref(sp), int(nsteps), C.byref(moved)))
        else:       # only the scheduled agents run schelling_step!; the others keep mood and position
            a = np.ascontiguousarray(ids, dtype=np.int64)
            model._ck(L.lib().ab_step_schelling_active(model.h, C.byref(sp), int(nsteps), model._ptr(a), a.shape[0], C.byref(moved)))
        self.last_moved = moved.value


class StandardABM:
    """StandardABM(space; model_step!, properties) with a device-resident SoA container.

    `model_step` is a rule object (FlockRule, PairForceRule, SchellingRule) — the
    B200Step of INTEGRATION.md; `step(n)` is `step!(model, n)`
    (src/simulations/step_standard.jl:24-31)."""

    def __init__(self, space, model_step=None, properties=None, device=0):
        self.space = space
        self.model_step = model_step
        self.properties = properties or {}
        self.time = 0
        self.lib = L.lib()
        desc = space._desc()
        h = C.c_void_p()
        rc = self.lib.ab_create(C.byref(desc), int(device), C.byref(h))
        if rc != L.AB_OK:
            raise AgentsError(rc, self.lib.ab_last_error(None).decode())
        self.h = h
        space.dims = tuple(int(desc.grid_dims[i]) for i in range(space.D))
        self.D = space.D
        self._columns = {"pos": 0}
        if space.kind == L.SPACE_CONTINUOUS:
            self._columns["vel"] = 1
        self._dtypes = {"pos": space.dtype, "vel": space.dtype}
        self._ncomp = {"pos": space.D, "vel": space.D}

    # -- plumbing ---------------------------------------------------------
    def close(self):
        if getattr(self, "h", None):
            self.lib.ab_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc < 0:
            raise AgentsError(rc, self.lib.ab_last_error(self.h).decode())
        return rc

    @staticmethod
    def _ptr(a):
        return a.ctypes.data_as(C.c_void_p) if a is not None else None

    # -- 1-D spaces embedded in 2-D, Float16 columns held as Float32 (ContinuousSpace docstring) --------------
    def _rows_in(self, a, fill=0.0):
        """caller's (n, user_D) rows -> the device's (n, D) rows of the device float type."""
        sp = self.space
        ud = getattr(sp, "user_D", self.D)
        a = np.asarray(a, dtype=sp.dtype).reshape(-1, ud)
        if ud == self.D:
            return np.ascontiguousarray(a)
        out = np.full((a.shape[0], self.D), fill, dtype=sp.dtype)
        out[:, :ud] = a
        return out

    def _rows_out(self, a):
        sp = self.space
        ud = getattr(sp, "user_D", self.D)
        if ud != self.D:
            a = a[:, :ud]
        udt = getattr(sp, "user_dtype", a.dtype)
        return a.astype(udt) if udt != a.dtype else a

    # -- population (add_agent! at setup; src/core/space_interaction_API.jl:174-179) ----
    def nagents(self):
        return int(self.lib.ab_nagents(self.h))

    def field_id(self, name):
        if name not in self._columns:
            raise KeyError(name)
        return self._columns[name]

    def add_agents(self, pos, vel=None, **columns):
        """Bulk add_agent!: ids 1..n in the given order; replaces the population."""
        dt = self.space.dtype
        if self.space.kind == L.SPACE_CONTINUOUS:
            pos = self._rows_in(pos, fill=0.5 * self.space.spacing)
        else:
            pos = np.ascontiguousarray(pos, dtype=dt).reshape(-1, self.D)
        n = pos.shape[0]
        self._ck(self.lib.ab_resize(self.h, n))
        self._ck(self.lib.ab_upload(self.h, 0, self._ptr(pos), n))
        if "vel" in self._columns:
            v = np.zeros((n, self.D), dtype=dt) if vel is None else self._rows_in(vel)
            self._ck(self.lib.ab_upload(self.h, 1, self._ptr(v), n))
        for name, col in columns.items():
            self.add_column(name, col)
        self._ck(self.lib.ab_build_index(self.h))
        return n

    def add_column(self, name, values):
        """One device column per isbits field of the @agent struct (src/core/agents.jl:199-231)."""
        a = np.ascontiguousarray(values)
        if a.dtype == np.bool_:
            a = a.astype(np.uint8)
        n = self.nagents()
        a = a.reshape(n, -1)
        if name not in self._columns:
            fid = self._ck(self.lib.ab_register_field(self.h, name.encode(), _NP2AB[a.dtype], a.shape[1]))
            self._columns[name] = fid
            self._dtypes[name] = a.dtype
            self._ncomp[name] = a.shape[1]
        a = np.ascontiguousarray(a, dtype=self._dtypes[name])
        self._ck(self.lib.ab_upload(self.h, self._columns[name], self._ptr(a), n))
        return self._columns[name]

    def _is_spatial_column(self, name):
        return self.space.kind == L.SPACE_CONTINUOUS and name in ("pos", "vel")

    def set_column(self, name, values):
        n = self.nagents()
        if self._is_spatial_column(name):
            a = self._rows_in(values, fill=0.5 * self.space.spacing if name == "pos" else 0.0)
            if a.shape[0] != n:
                raise AgentsError(L.ERR_INVALID, f"set_column({name!r}): one row per agent")
        else:
            a = np.ascontiguousarray(values, dtype=self._dtypes[name]).reshape(n, self._ncomp[name])
        self._ck(self.lib.ab_upload(self.h, self._columns[name], self._ptr(a), n))

    def _column_raw(self, name):
        """the device column as stored (checkpoints: D device components, Float16 agents still in Float32)."""
        n = self.nagents()
        out = np.zeros((n, self._ncomp[name]), dtype=self._dtypes[name])
        self._ck(self.lib.ab_download(self.h, self._columns[name], self._ptr(out), n))
        return out if self._ncomp[name] > 1 else out[:, 0]

    def column(self, name):
        out = self._column_raw(name)
        if self._is_spatial_column(name):
            return self._rows_out(out)            # (n, user_D) of the caller's float type, also on an embedded 1-D space
        return out

    @property
    def pos(self):
        p = self.column("pos")
        return self._rows_out(p) if self.space.kind == L.SPACE_CONTINUOUS else p

    @property
    def vel(self):
        return self._rows_out(self.column("vel"))

    def set_refinement(self, K):
        """Device-grid refinement (1 or 2 sub-cells per reference cell and dimension); results do not depend on it."""
        self._ck(self.lib.ab_set_refinement(self.h, int(K)))

    def cells(self):
        """pos2cell of every agent, 1-based (continuous.jl:112-113)."""
        n = self.nagents()
        out = np.zeros((n, self.D), dtype=np.int32)
        self._ck(self.lib.ab_cells(self.h, self._ptr(out), n))
        return out[:, : getattr(self.space, "user_D", self.D)]

    # -- neighbour queries ---------------------------------------------------
    def nearby_ids_batch(self, positions=None, ids=None, r=1.0, search="approximate", exclude_self=True):
        """Batched nearby_ids.  Returns (offsets[nq+1], ids) CSR with 1-based ids."""
        if (positions is None) == (ids is None):
            raise ValueError("give positions or ids")
        if search not in L.SEARCH:
            raise AgentsError(L.ERR_INVALID, "`search` keyword should be either `:approximate` or `:exact`")
        if positions is not None:
            if self.space.kind == L.SPACE_CONTINUOUS:
                q = self._rows_in(positions, fill=0.5 * self.space.spacing)
            else:
                q = np.ascontiguousarray(positions, dtype=self.space.dtype).reshape(-1, self.D)
            nq, qp, qi = q.shape[0], self._ptr(q), None
        else:
            q = np.ascontiguousarray(ids, dtype=np.int64).reshape(-1)
            nq, qp, qi = q.shape[0], None, self._ptr(q)
        counts = np.zeros(nq, dtype=np.int64)
        self._ck(self.lib.ab_nearby_count(self.h, qp, qi, nq, float(r), L.SEARCH[search], int(exclude_self), self._ptr(counts)))
        offsets = np.zeros(nq + 1, dtype=np.int64)
        np.cumsum(counts, out=offsets[1:])
        out = np.zeros(max(int(offsets[-1]), 1), dtype=np.int64)
        self._ck(self.lib.ab_nearby_fill(self.h, qp, qi, nq, float(r), L.SEARCH[search], int(exclude_self),
                                         self._ptr(offsets), self._ptr(out)))
        return offsets, out[: int(offsets[-1])]

    def nearby_ids(self, x, r=1.0, search="approximate"):
        """nearby_ids(pos, model, r; search) for a position, nearby_ids(agent, ...) for an id
        (the agent form drops the agent's own id, space_interaction_API.jl:284-287)."""
        if isinstance(x, (int, np.integer)):
            _, ids = self.nearby_ids_batch(ids=[int(x)], r=r, search=search, exclude_self=True)
        else:
            _, ids = self.nearby_ids_batch(positions=[x], r=r, search=search)
        return ids

    def interacting_pairs(self, r, method="all", search="exact", scheduler=None):
        """interacting_pairs(model, r, method; search, scheduler) → (P, 2) array of id pairs
        (continuous.jl:418-436).  `:all` (all_pairs!, :438-452) and `:nearest` (true_pairs!, :454-497;
        `scheduler` = the id order, default ByID) run on the device; `:types` is not built."""
        if method == "nearest":
            order = None if scheduler is None else np.ascontiguousarray(scheduler, dtype=np.int64)
            npairs = C.c_int64(0)
            self._ck(self.lib.ab_pairs_nearest(self.h, float(r), self._ptr(order) if order is not None else None,
                                               0 if order is None else len(order), C.byref(npairs)))
            P = npairs.value
            lo = np.zeros(max(P, 1), dtype=np.int64)
            hi = np.zeros(max(P, 1), dtype=np.int64)
            self._ck(self.lib.ab_pairs_nearest_fill(self.h, self._ptr(lo), self._ptr(hi), P))
            return np.stack([lo[:P], hi[:P]], axis=1)
        if method != "all":
            raise NotImplementedError("method = :types is not built (agent unions are out of scope)")
        npairs = C.c_int64(0)
        self._ck(self.lib.ab_pairs_count(self.h, float(r), L.SEARCH[search], C.byref(npairs)))
        P = npairs.value
        lo = np.zeros(max(P, 1), dtype=np.int64)
        hi = np.zeros(max(P, 1), dtype=np.int64)
        self._ck(self.lib.ab_pairs_fill(self.h, self._ptr(lo), self._ptr(hi), P))
        return np.stack([lo[:P], hi[:P]], axis=1)

    def elastic_collisions(self, pairs, mass=None):
        """elastic_collision!(a, b, f) for every pair of `pairs` in order (continuous.jl:333-366);
        `mass` = name of the mass column (the `f` symbol) or None.  Returns the per-pair `happened` flags."""
        pairs = np.ascontiguousarray(pairs, dtype=np.int64).reshape(-1, 2)
        a, b = np.ascontiguousarray(pairs[:, 0]), np.ascontiguousarray(pairs[:, 1])
        hap = np.zeros(max(len(a), 1), dtype=np.uint8)
        nc = C.c_int64(0)
        self._ck(self.lib.ab_elastic_collisions(self.h, self._ptr(a), self._ptr(b), len(a),
                                                -1 if mass is None else self._columns[mass], self._ptr(hap), C.byref(nc)))
        return hap[: len(a)].astype(bool)

    def nearest_neighbor(self, r):
        """nearest_neighbor(agent, model, r) for every agent: id or 0 for `nothing` (continuous.jl:298-307)."""
        n = self.nagents()
        out = np.zeros(n, dtype=np.int64)
        self._ck(self.lib.ab_nearest_neighbor(self.h, float(r), self._ptr(out), n))
        return out

    def nearby_positions_batch(self, positions, r=1, empty=False):
        """nearby_positions(pos, model, r) (grid_general.jl:185-232) / empty_nearby_positions (discrete.jl:116-129)
        for a batch of grid positions: CSR (offsets, (total, D) positions)."""
        q = np.ascontiguousarray(positions, dtype=np.int32).reshape(-1, self.D)
        nq = len(q)
        counts = np.zeros(max(nq, 1), dtype=np.int64)
        self._ck(self.lib.ab_nearby_positions_count(self.h, self._ptr(q), nq, float(r), int(bool(empty)), self._ptr(counts)))
        off = np.zeros(nq + 1, dtype=np.int64)
        np.cumsum(counts[:nq], out=off[1:])
        out = np.zeros((max(int(off[-1]), 1), self.D), dtype=np.int32)
        self._ck(self.lib.ab_nearby_positions_fill(self.h, self._ptr(q), nq, float(r), int(bool(empty)), self._ptr(off), self._ptr(out)))
        return off, out[: int(off[-1])]

    def nearby_positions(self, pos, r=1):
        return self.nearby_positions_batch([pos], r)[1]

    def empty_nearby_positions(self, pos, r=1):
        return self.nearby_positions_batch([pos], r, empty=True)[1]

    # -- spatial fields ---------------------------------------------------------
    def _field_args(self, prop):
        prop = np.asarray(prop, dtype=np.float64)
        return prop, self._field_dims(prop.shape)

    def _field_dims(self, shape):
        """dimensions of a spatial property: at most the space's own D (1 on an embedded 1-D space)."""
        dims = np.ascontiguousarray(shape, dtype=np.int64).reshape(-1)
        ud = getattr(self.space, "user_D", self.D)
        if not 1 <= len(dims) <= ud:
            raise AgentsError(L.ERR_INVALID, f"a spatial property of a {ud}-D space has 1..{ud} dimensions, got {len(dims)}")
        return dims

    def _query_rows(self, positions):
        if positions is None:
            return None
        if self.space.kind == L.SPACE_CONTINUOUS:
            return self._rows_in(positions, fill=0.5 * self.space.spacing)
        return np.ascontiguousarray(positions, dtype=self.space.dtype).reshape(-1, self.D)

    def get_spatial_index(self, shape, positions=None):
        """get_spatial_index(pos, property, model) (continuous.jl:566-573) for all agents (id order) or for
        `positions`: (n, len(shape)) 1-based indices into an array of `shape`."""
        dims = self._field_dims(shape)
        q = self._query_rows(positions)
        n = self.nagents() if q is None else len(q)
        out = np.zeros((max(n, 1), len(dims)), dtype=np.int64)
        self._ck(self.lib.ab_spatial_lookup(self.h, self._ptr(dims), len(dims), None, 0, self._ptr(q) if q is not None else None,
                                            n, self._ptr(out), None))
        return out[:n]

    def get_spatial_property(self, prop, positions=None):
        """get_spatial_property(pos, property::AbstractArray, model) (continuous.jl:545-548), batched."""
        prop, dims = self._field_args(prop)
        flat = np.ascontiguousarray(prop.ravel(order="F"))
        q = self._query_rows(positions)
        n = self.nagents() if q is None else len(q)
        out = np.zeros(max(n, 1), dtype=np.float64)
        self._ck(self.lib.ab_spatial_lookup(self.h, self._ptr(dims), len(dims), self._ptr(flat), 1,
                                            self._ptr(q) if q is not None else None, n, None, self._ptr(out)))
        return out[:n]

    def update_vel_field(self, field, scale=1.0, ids=None):
        """The force-field `update_vel!` hook (continuous.jl:86-91): vel += scale * field[index(pos)], where
        `field` has shape property_shape + (D,)."""
        field = np.asarray(field, dtype=np.float64)
        ud = getattr(self.space, "user_D", self.D)
        if field.ndim < 2 or field.shape[-1] != ud:
            raise AgentsError(L.ERR_INVALID, f"update_vel!: the field's last axis holds the {ud} velocity components")
        dims = self._field_dims(field.shape[:-1])
        comps = [field[..., d].ravel(order="F") for d in range(ud)]
        comps += [np.zeros_like(comps[0])] * (self.D - ud)        # embedded 1-D space: no force across the row
        flat = np.ascontiguousarray(np.concatenate(comps))
        a = np.ascontiguousarray(ids, dtype=np.int64) if ids is not None else None
        self._ck(self.lib.ab_update_vel_field(self.h, self._ptr(dims), len(dims), self._ptr(flat), float(scale),
                                              self._ptr(a) if a is not None else None, 0 if a is None else len(a)))
        return self

    # -- movement --------------------------------------------------------------
    def move_agent(self, dt, ids=None):
        """move_agent!(agent, model, dt) for `ids` (all agents when None) — continuous.jl:186-194."""
        a = np.ascontiguousarray(ids, dtype=np.int64) if ids is not None else None
        self._ck(self.lib.ab_move(self.h, float(dt), self._ptr(a), 0 if a is None else a.shape[0]))

    def walk(self, directions, ids=None, ifempty=True):
        """walk!(agent, direction, model) — walk.jl:33-66."""
        dt = self.space.dtype
        if self.space.kind == L.SPACE_CONTINUOUS:
            d = self._rows_in(directions)
        else:
            d = np.ascontiguousarray(directions, dtype=dt).reshape(-1, self.D)
        a = np.ascontiguousarray(ids, dtype=np.int64) if ids is not None else None
        if d.shape[0] != (self.nagents() if a is None else a.shape[0]):
            raise AgentsError(L.ERR_INVALID, "walk!: one direction per walking agent")
        self._ck(self.lib.ab_walk(self.h, self._ptr(d), self._ptr(a), d.shape[0] if a is None else a.shape[0], int(ifempty)))

    def randomwalk_angles(self, polar, azimuthal=None, r=None, ids=None):
        """randomwalk!(agent, model[, r]; polar, azimuthal) on a ContinuousSpace (walk.jl:236-330): `polar` /
        `azimuthal` are the angles drawn for the walking agents (arrays, or callables k -> k angles, e.g. a
        distribution's sampler); the velocity is rotated by them, scaled to length r when r is given, and walked."""
        if getattr(self.space, "user_D", self.D) == 1:
            raise AgentsError(L.ERR_INVALID, "randomwalk! with angles needs a 2-D or 3-D ContinuousSpace (walk.jl:236-330)")
        n = self.nagents()
        a = np.ascontiguousarray(ids, dtype=np.int64) if ids is not None else None
        k = n if a is None else a.shape[0]
        th = np.ascontiguousarray(polar(k) if callable(polar) else polar, dtype=np.float64).reshape(k)
        ph = None
        if self.D == 3:
            if azimuthal is None:
                raise AgentsError(-1, "3-D random walks need azimuthal angles")
            ph = np.ascontiguousarray(azimuthal(k) if callable(azimuthal) else azimuthal, dtype=np.float64).reshape(k)
        if r is not None and r <= 0:
            raise AgentsError(-1, "ArgumentError: The displacement must be larger than 0.")
        self._ck(self.lib.ab_randomwalk_angles(self.h, -1.0 if r is None else float(r), self._ptr(th), self._ptr(ph),
                                               self._ptr(a), 0 if a is None else k))

    def randomwalk(self, r=None, seed=0, ids=None, ifempty=True, force_motion=False):
        """randomwalk!(agent, model, r; ifempty, force_motion) for `ids` (all agents when None) — walk.jl:154-214 on
        grids (r defaults to 1), walk.jl:389-408 on a ContinuousSpace (r defaults to the agent's speed).  `seed`
        selects the counter stream of this call (see ab_randomwalk in include/agents_b200.h)."""
        if getattr(self.space, "user_D", self.D) == 1:
            raise AgentsError(L.ERR_INVALID, "randomwalk! on an embedded 1-D ContinuousSpace is not built (the device draws D = 2 normals)")
        if r is None:
            r = -1.0 if isinstance(self.space, ContinuousSpace) else 1.0
        elif isinstance(self.space, ContinuousSpace) and r <= 0:
            raise AgentsError(-1, "ArgumentError: The displacement must be larger than 0.")
        a = np.ascontiguousarray(ids, dtype=np.int64) if ids is not None else None
        self._ck(self.lib.ab_randomwalk(self.h, float(r), int(seed) & 0xFFFFFFFFFFFFFFFF, self._ptr(a),
                                        0 if a is None else a.shape[0], int(ifempty), int(force_motion)))

    # -- stepping (src/simulations/step_standard.jl:24-31) --------------------
    def step(self, n=1, scheduler=None, check_every=1):
        """step!(model, n) or step!(model, f) (src/simulations/step.jl:7-27, step_standard.jl:9-31).

        `n`: an integer — exactly n ticks, batched into one ccall — or a predicate f(model, t): the model steps until
        f(model, t) returns true, t = ticks taken by this call (`until(t1, t0, f, model) = !f(model, t1 - t0)`,
        step.jl:38).  The predicate runs on the host between device calls of `check_every` ticks (1 = the reference's
        tick-by-tick evaluation).
        `scheduler`: None (every agent: fastest / ByID / Randomly / ByProperty are equivalent under the synchronous
        update), an id list, or a callable model -> id list evaluated every tick (Schedulers.Partially and custom
        schedulers, schedulers.jl:8, 68-120); a scheduled subset steps, the other agents keep their state."""
        if self.model_step is None:
            raise AgentsError(L.ERR_STATE, "no model_step! rule")
        if callable(n):
            t0 = self.time
            while not n(self, self.time - t0):
                if scheduler is not None:
                    self.step(1, scheduler=scheduler)
                else:
                    self.model_step(self, int(check_every))
                    self.time += int(check_every)
            return self
        if scheduler is not None:
            for _ in range(int(n)):
                self.model_step(self, 1, ids=scheduler(self) if callable(scheduler) else scheduler)
                self.time += 1
            return self
        self.model_step(self, int(n))      # n ticks batched into one ccall
        self.time += int(n)
        return self

    # -- schedulers (src/submodules/schedulers.jl) ------------------------------
    def schedule_by_property(self, field, comp=0):
        """Schedulers.ByProperty(field)(model) — schedulers.jl:130-151: the agent ids ordered by `sortperm` of the
        property column (ascending, stable: Julia's default sort algorithm for sortperm is stable).  Sorted on the
        device (ab_ids_by_property); only the ids come back."""
        n = self.nagents()
        out = np.zeros(n, dtype=np.int64)
        self._ck(self.lib.ab_ids_by_property(self.h, self.field_id(field), int(comp), 0, self._ptr(out), n))
        return out

    def schedule_partially(self, p, rng):
        """Schedulers.Partially(p)(model) — schedulers.jl:109-120: randsubseq of the ids with probability p (each id is
        kept independently; the reference draws from abmrng(model), whose stream is not reproduced: pass a numpy Generator)."""
        n = self.nagents()
        return (np.nonzero(rng.random(n) < p)[0] + 1).astype(np.int64)

    # -- data collection (src/simulations/collect.jl:154-183, 466-502) ----------
    def collect(self, adata):
        """collect_agent_data! for aggregate specs: adata = [(field, aggregator[, filter])], field = "name" or
        ("name", component); aggregator in sum / mean / minimum / maximum / count; filter = (field, cmp, value)
        with cmp in == != < <= > >=.  Reduced on the device; returns one float per spec."""
        specs = (L.ReduceSpec * len(adata))()
        for i, item in enumerate(adata):
            field, agg = item[0], item[1]
            flt = item[2] if len(item) > 2 else None
            name, comp = (field, 0) if isinstance(field, str) else field
            agg = getattr(agg, "__name__", agg)
            if agg not in L.RED:
                raise ValueError(f"aggregator {agg!r} does not run on the device")
            specs[i].field, specs[i].comp, specs[i].op = self.field_id(name), int(comp), L.RED[agg]
            specs[i].filter_cmp = 0
            if flt is not None:
                ff, cmp_, val = flt
                fname, fcomp = (ff, 0) if isinstance(ff, str) else ff
                specs[i].filter_field, specs[i].filter_comp = self.field_id(fname), int(fcomp)
                specs[i].filter_cmp, specs[i].filter_value = L.CMP[cmp_], float(val)
        out = np.zeros(len(adata), dtype=np.float64)
        self._ck(self.lib.ab_reduce(self.h, specs, len(adata), self._ptr(out), None))
        return out

    def run(self, n, adata=None, when=1):
        """run!(model, n; adata, when): steps n ticks, collecting the aggregates of `adata` at time 0 and after
        every `when`-th tick (collect.jl:154-183).  Returns {"time": [...], "<agg>_<field>": [...]} —
        the columns of the reference's agent DataFrame for tuple-form adata."""
        adata = list(adata or [])

        def label(item):
            field, agg = item[0], item[1]
            name = field if isinstance(field, str) else f"{field[0]}[{field[1] + 1}]"
            return f"{getattr(agg, '__name__', agg)}_{name}" + ("_filtered" if len(item) > 2 else "")

        cols = {"time": []}
        labels = [label(it) for it in adata]
        for lb in labels:
            cols[lb] = []

        def take():
            cols["time"].append(self.time)
            if adata:
                for lb, v in zip(labels, self.collect(adata)):
                    cols[lb].append(float(v))

        take()
        done = 0
        while done < n:
            k = min(int(when), n - done)
            self.step(k)                       # `when` ticks batched into one call
            done += k
            if done % int(when) == 0:
                take()
        return cols

    # -- checkpoints (AgentsIO, src/submodules/io/jld2_integration.jl:85-135, 196-223) ----------
    def save_checkpoint(self, path):
        """AgentsIO.save_checkpoint(filename, model): every column in id order, the space, the rule, the model
        properties, the time, and the device layout a bit-for-bit continuation needs (ab_layout_get: the order of
        the agents inside the cells — the reference's stored_ids — the grid refinement, the relocation counter).
        Written as one .npz; properties must be JSON-serialisable, as JLD2 needs them serialisable."""
        import json
        n = self.nagents()
        sp = self.space
        rule = self.model_step
        meta = {"format": "agents_b200-checkpoint-1", "time": int(self.time), "n": n, "properties": self.properties,
                "space": {"kind": type(sp).__name__, "periodic": [bool(x) for x in sp.periodic], "metric": sp.metric},
                "columns": [[name, str(self._dtypes[name]), int(self._ncomp[name])]
                            for name, _ in sorted(self._columns.items(), key=lambda kv: kv[1])],
                "rule": None}
        if isinstance(sp, ContinuousSpace):
            meta["space"].update(extent=list(sp.extent), spacing=sp.spacing, dtype=str(sp.dtype))
        else:
            meta["space"].update(dims=[int(x) for x in sp.dims])
        arrays = {}
        if isinstance(rule, FlockRule):
            meta["rule"] = {"type": "FlockRule", "p": rule.p, "search": rule.search, "kernel": rule.kernel,
                            "field": rule._field}
            if rule.per_agent is not None and rule._field < 0:
                arrays["rule_per_agent"] = np.asarray(rule.per_agent)
        elif isinstance(rule, PairForceRule):
            meta["rule"] = {"type": "PairForceRule", "dt": rule.dt, "cutoff": rule.cutoff}
        elif isinstance(rule, SchellingRule):
            meta["rule"] = {"type": "SchellingRule", "min_to_be_happy": rule.min_to_be_happy,
                            "relocate": bool(rule.relocate), "seed": int(rule.seed)}
        elif rule is not None:
            raise TypeError("only the device rules can be checkpointed")
        for name in self._columns:
            arrays["col_" + name] = self._column_raw(name)
        order = np.zeros(n, dtype=np.int64)
        K, tick = C.c_int32(1), C.c_int64(0)
        self._ck(self.lib.ab_layout_get(self.h, self._ptr(order), n, C.byref(K), C.byref(tick)))
        meta["refinement"], meta["tick"] = int(K.value), int(tick.value)
        arrays["slot_order"] = order
        with open(path, "wb") as f:
            np.savez(f, meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8), **arrays)

    def last_neighbor_counts(self):
        n = self.nagents()
        out = np.zeros(n, dtype=np.int32)
        self._ck(self.lib.ab_flock_last_counts(self.h, self._ptr(out), n))
        return out

    def set_async(self, on=True):
        """Asynchronous mode (ab_set_async): upload / step / download calls return once queued on the
        model's stream; `sync()` waits and raises what failed on the device in between.  Buffers handed
        to `upload_rows` / `download_rows` must be pinned and stay alive until `sync()`."""
        self._ck(self.lib.ab_set_async(self.h, 1 if on else 0))
        return self

    def sync(self):
        self._ck(self.lib.ab_sync(self.h))
        return self

    def upload_rows(self, name, rows):
        """Column `name` from a caller-owned (ideally pinned) row-per-agent array, id order; no copy is made."""
        self._ck(self.lib.ab_upload(self.h, self._columns[name], self._ptr(rows), self.nagents()))
        return self

    def download_rows(self, name, out):
        """Column `name` into a caller-owned (ideally pinned) row-per-agent array, id order."""
        self._ck(self.lib.ab_download(self.h, self._columns[name], self._ptr(out), self.nagents()))
        return out

    def launch_count(self):
        return int(self.lib.ab_launch_count(self.h))

    def last_step_ms(self):
        return float(self.lib.ab_last_step_ms(self.h))


def load_checkpoint(path, device=0):
    """AgentsIO.load_checkpoint(filename) (jld2_integration.jl:196-223): a StandardABM on `device` whose state, rule,
    properties and time equal the saved model's, and whose continuation equals the uninterrupted run bit for bit."""
    import json
    with np.load(path, allow_pickle=False) as z:
        meta = json.loads(bytes(z["meta"]).decode())
        if meta.get("format") != "agents_b200-checkpoint-1":
            raise ValueError("not a checkpoint of this library")
        arrays = {k: z[k] for k in z.files if k != "meta"}
    sm = meta["space"]
    if sm["kind"] == "ContinuousSpace":
        space = ContinuousSpace(sm["extent"], spacing=sm["spacing"], periodic=tuple(sm["periodic"]), dtype=np.dtype(sm["dtype"]))
    else:
        space = {"GridSpace": GridSpace, "GridSpaceSingle": GridSpaceSingle}[sm["kind"]](
            tuple(sm["dims"]), periodic=tuple(sm["periodic"]), metric=sm["metric"])
    rm, rule = meta["rule"], None
    if rm is not None:
        if rm["type"] == "FlockRule":
            kern = {0: "auto", 1: "generic", 2: "tiled"}[rm["kernel"]]
            rule = FlockRule(search=rm["search"], kernel=kern, per_agent=arrays.get("rule_per_agent"), **rm["p"])
        elif rm["type"] == "PairForceRule":
            rule = PairForceRule(dt=rm["dt"], cutoff=rm["cutoff"])
        else:
            rule = SchellingRule(min_to_be_happy=rm["min_to_be_happy"], relocate=rm["relocate"], seed=rm["seed"])
    m = StandardABM(space, model_step=rule, properties=meta["properties"], device=device)
    n = meta["n"]
    vel = arrays.get("col_vel")
    m.add_agents(arrays["col_pos"].reshape(n, space.D), None if vel is None else vel.reshape(n, space.D))
    for name, _, _ in meta["columns"]:                       # registration order = field ids of the saved model
        if name not in ("pos", "vel"):
            m.add_column(name, arrays["col_" + name])
    if rm is not None and rm["type"] == "FlockRule" and rm["field"] >= 0:
        rule._field = m.field_id("_flock_params")
        rule.per_agent = arrays["col__flock_params"]
    m._ck(m.lib.ab_layout_set(m.h, m._ptr(np.ascontiguousarray(arrays["slot_order"], dtype=np.int64)), n,
                              int(meta["refinement"]), int(meta["tick"])))
    m.time = int(meta["time"])
    return m
