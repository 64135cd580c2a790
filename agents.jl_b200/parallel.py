"""Host side of the multi-GPU slab decomposition (one process per GPU).

The reference is single-task (src/simulations/step_standard.jl:9-23); this module is the host
plumbing of its scale-out: which cell layers of the last axis a rank owns, which agents
it must hold (owned + halo layers), and how a global id-ordered state is scattered to / gathered
from the ranks.  `torch.distributed` (NCCL on GPUs, gloo in the CPU tests) only moves the 128-byte
communicator id and the gathered results; the per-step halo / migration exchange happens inside the
library (csrc/ab_comm.cu).

Pure-numpy layout functions (`slab_layout`, `window_layers`, `cell_layer`, `select_window`) carry no
device dependency so they are tested on CPU with world_size > 1.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

HALO_W = 3   # layers per side; must equal AB_HALO_W (csrc/ab_comm.cu) — checked against ab_comm_halo()


def slab_layout(n_layers: int, nranks: int):
    """[(first_layer, count)] per rank: the first (n_layers % nranks) ranks own one layer more."""
    base, rem = divmod(int(n_layers), int(nranks))
    out = []
    for r in range(nranks):
        out.append((r * base + min(r, rem), base + (1 if r < rem else 0)))
    return out


def window_layers(first: int, count: int, n_layers: int, W: int = HALO_W, periodic: bool = True):
    """Global layers a rank stores, in its local order: owned, upper halo, lower halo.  On a non-periodic axis the
    halo layers beyond the space do not exist (None)."""
    own = [(first + i) % n_layers for i in range(count)]
    up = [(first + count + i) % n_layers if periodic or first + count + i < n_layers else None for i in range(W)]
    lo = [(first - W + i) % n_layers if periodic or first - W + i >= 0 else None for i in range(W)]
    return own, up, lo


def cell_layer(pos_last, spacing, n_layers):
    """0-based cell layer of the last coordinate: max(1, ceil(p / spacing)) - 1 (continuous.jl:112-113),
    IEEE division in the column's own precision — the same arithmetic as the device's pos2cell."""
    p = np.asarray(pos_last)
    c = np.ceil(p / p.dtype.type(spacing)).astype(np.int64)
    return np.clip(np.maximum(c, 1) - 1, 0, n_layers - 1)


def select_window(layer, first, count, n_layers, W: int = HALO_W, periodic: bool = True):
    """Indices (into the global arrays) of the agents a rank must hold, and the owned subset mask."""
    if periodic:
        rel = (np.asarray(layer, dtype=np.int64) - first) % n_layers      # 0 .. n_layers-1
        owned = rel < count
        halo = ((rel >= count) & (rel < count + W)) | (rel >= n_layers - W)
    else:                                                                 # no wrap: the end slabs have one halo only
        rel = np.asarray(layer, dtype=np.int64) - first
        owned = (rel >= 0) & (rel < count)
        halo = ((rel >= count) & (rel < count + W)) | ((rel < 0) & (rel >= -W))
    idx = np.nonzero(owned | halo)[0]
    return idx, owned[idx]


def default_halo_capacity(n_global: int, n_layers: int, W: int = HALO_W, factor: float = 3.0) -> int:
    """Message capacity every rank agrees on: `factor` x the mean population of W layers + slack."""
    return int(factor * W * (n_global / max(n_layers, 1))) + 4096


class Slabs:
    """Sharded view of a StandardABM over `torch.distributed` ranks."""

    def __init__(self, model, dist=None, rank=None, world=None, init_nccl=True):
        from . import _lib as L

        self.model, self.dist, self.L = model, dist, L
        self.init_nccl = init_nccl
        self.rank = int(dist.get_rank() if rank is None else rank)
        self.world = int(dist.get_world_size() if world is None else world)
        lib = model.lib
        uid = None
        if init_nccl:
            buf = (C.c_ubyte * 128)()
            if self.rank == 0:
                model._ck(lib.ab_comm_unique_id(buf))
            obj = [bytes(buf) if self.rank == 0 else None]
            dist.broadcast_object_list(obj, src=0)
            uid = (C.c_ubyte * 128).from_buffer_copy(obj[0])
        model._ck(lib.ab_comm_init(model.h, self.rank, self.world, uid))
        first, count = C.c_int64(0), C.c_int64(0)
        model._ck(lib.ab_comm_slab(model.h, C.byref(first), C.byref(count)))
        self.first, self.count = first.value, count.value
        self.n_layers = model.space.dims[-1]
        self.periodic_last = bool(model.space.periodic[-1])
        assert (self.first, self.count) == slab_layout(self.n_layers, self.world)[self.rank]
        assert lib.ab_comm_halo(model.h) == HALO_W
        self.n_global = 0

    def _link_direct(self):
        """Direct halo exchange over peer memory (ab_comm_ipc_*): every rank exports the CUDA IPC handle of
        its receive block, the handles go round through torch.distributed, every rank maps its two
        neighbours' blocks.  AB_HALO=nccl keeps the ncclSend/ncclRecv exchange (the baseline)."""
        import os

        m, lib = self.model, self.model.lib
        if self.dist is None or not self.init_nccl or os.environ.get("AB_HALO", "direct") == "nccl":
            return False
        mine = (C.c_ubyte * 64)()
        rc = lib.ab_comm_ipc_export(m.h, mine)
        handles = [None] * self.world
        self.dist.all_gather_object(handles, bytes(mine) if rc == 0 else None)
        if any(h is None for h in handles):
            return False                                    # some rank could not export: everyone stays on NCCL
        has_lo = self.periodic_last or self.rank > 0                 # end slabs of a non-periodic axis have one neighbour
        has_up = self.periodic_last or self.rank < self.world - 1
        lo = (C.c_ubyte * 64).from_buffer_copy(handles[(self.rank - 1) % self.world]) if has_lo else None
        up = (C.c_ubyte * 64).from_buffer_copy(handles[(self.rank + 1) % self.world]) if has_up else None
        rc = lib.ab_comm_ipc_import(m.h, lo, up)
        oks = [None] * self.world
        self.dist.all_gather_object(oks, rc == 0)
        if not all(oks):
            raise RuntimeError("direct halo exchange: a rank failed to map its neighbours (set AB_HALO=nccl)")
        return True

    # -- population -----------------------------------------------------------------------------
    def scatter(self, pos, vel, halo_capacity=None, capacity_factor=1.6, columns=None):
        """Every rank passes the same global, id-ordered (1..N) state and keeps its window of it.  `columns`:
        {name: global id-ordered array} of user columns (the r / k / mass of the pair-force rule, ...); they are
        registered BEFORE the receive block is exported, because they travel in the boundary messages."""
        from .model import _NP2AB

        m, lib = self.model, self.model.lib
        dt = m.space.dtype
        pos = np.ascontiguousarray(pos, dtype=dt).reshape(-1, m.D)
        vel = np.ascontiguousarray(vel, dtype=dt).reshape(-1, m.D)
        self.n_global = pos.shape[0]
        cols = {}
        for name, a in (columns or {}).items():
            a = np.ascontiguousarray(a).reshape(self.n_global, -1)
            if name not in m._columns:
                fid = m._ck(lib.ab_register_field(m.h, name.encode(), _NP2AB[a.dtype], a.shape[1]))
                m._columns[name], m._dtypes[name], m._ncomp[name] = fid, a.dtype, a.shape[1]
            cols[name] = np.ascontiguousarray(a, dtype=m._dtypes[name])
        layer = cell_layer(pos[:, -1], m.space.spacing, self.n_layers)
        idx, owned = select_window(layer, self.first, self.count, self.n_layers, periodic=self.periodic_last)
        cap = halo_capacity or default_halo_capacity(self.n_global, self.n_layers)
        m._ck(lib.ab_comm_set_capacity(m.h, int(cap)))
        self.direct = self._link_direct()
        m._ck(lib.ab_resize(m.h, int(len(idx) * capacity_factor) + 2 * int(cap) + 1024))
        gids = (idx + 1).astype(np.int64)
        p = np.ascontiguousarray(pos[idx])
        v = np.ascontiguousarray(vel[idx])
        m._ck(lib.ab_upload_local(m.h, 0, m._ptr(p), m._ptr(gids), len(idx)))
        m._ck(lib.ab_upload_local(m.h, 1, m._ptr(v), m._ptr(gids), len(idx)))
        for name, a in cols.items():
            c = np.ascontiguousarray(a[idx])
            m._ck(lib.ab_upload_local(m.h, m._columns[name], m._ptr(c), m._ptr(gids), len(idx)))
        m._ck(lib.ab_build_index(m.h))
        return int(owned.sum())

    def local_column(self, name):
        """(global ids, values) of a user column for the agents this rank owns now."""
        m, lib = self.model, self.model.lib
        cap = int(lib.ab_nagents(m.h)) + 1
        ids = np.zeros(cap, dtype=np.int64)
        out = np.zeros((cap, m._ncomp[name]), dtype=m._dtypes[name])
        n = C.c_int64(0)
        m._ck(lib.ab_download_local(m.h, m._columns[name], m._ptr(out), m._ptr(ids), cap, C.byref(n)))
        return ids[: n.value], out[: n.value]

    def scatter_chunks(self, chunks, n_global, halo_capacity=None, capacity_factor=1.6):
        """Like `scatter` for populations too large to materialise on every rank: `chunks` yields
        (first_id0, pos_chunk, vel_chunk[, {column: chunk}]) pieces of the global id-ordered state; only this rank's
        window of each piece is kept.  User columns are registered before the receive block is exported."""
        from .model import _NP2AB

        m, lib = self.model, self.model.lib
        dt = m.space.dtype
        self.n_global = int(n_global)
        keep_ids, keep_p, keep_v, keep_c, owned_n = [], [], [], {}, 0
        for piece in chunks:
            first, pos, vel = piece[:3]
            cols = piece[3] if len(piece) > 3 else {}
            pos = np.ascontiguousarray(pos, dtype=dt).reshape(-1, m.D)
            vel = np.ascontiguousarray(vel, dtype=dt).reshape(-1, m.D)
            layer = cell_layer(pos[:, -1], m.space.spacing, self.n_layers)
            idx, owned = select_window(layer, self.first, self.count, self.n_layers, periodic=self.periodic_last)
            keep_ids.append((idx + first + 1).astype(np.int64))
            keep_p.append(pos[idx])
            keep_v.append(vel[idx])
            for name, a in cols.items():
                keep_c.setdefault(name, []).append(np.asarray(a).reshape(pos.shape[0], -1)[idx])
            owned_n += int(owned.sum())
        gids = np.concatenate(keep_ids)
        p = np.ascontiguousarray(np.concatenate(keep_p))
        v = np.ascontiguousarray(np.concatenate(keep_v))
        cols = {name: np.concatenate(parts) for name, parts in keep_c.items()}
        for name, a in cols.items():
            if name not in m._columns:
                fid = m._ck(lib.ab_register_field(m.h, name.encode(), _NP2AB[a.dtype], a.shape[1]))
                m._columns[name], m._dtypes[name], m._ncomp[name] = fid, a.dtype, a.shape[1]
        cap = halo_capacity or default_halo_capacity(self.n_global, self.n_layers)
        m._ck(lib.ab_comm_set_capacity(m.h, int(cap)))
        self.direct = self._link_direct()
        m._ck(lib.ab_resize(m.h, int(len(gids) * capacity_factor) + 2 * int(cap) + 1024))
        m._ck(lib.ab_upload_local(m.h, 0, m._ptr(p), m._ptr(gids), len(gids)))
        m._ck(lib.ab_upload_local(m.h, 1, m._ptr(v), m._ptr(gids), len(gids)))
        for name, a in cols.items():
            c = np.ascontiguousarray(a, dtype=m._dtypes[name])
            m._ck(lib.ab_upload_local(m.h, m._columns[name], m._ptr(c), m._ptr(gids), len(gids)))
        m._ck(lib.ab_build_index(m.h))
        return owned_n

    def local_state(self):
        """(global ids, pos, vel) of the agents this rank owns now."""
        m, lib = self.model, self.model.lib
        cap = int(lib.ab_nagents(m.h)) + 1
        ids = np.zeros(cap, dtype=np.int64)
        pos = np.zeros((cap, m.D), dtype=m.space.dtype)
        vel = np.zeros((cap, m.D), dtype=m.space.dtype)
        n = C.c_int64(0)
        m._ck(lib.ab_download_local(m.h, 0, m._ptr(pos), m._ptr(ids), cap, C.byref(n)))
        m._ck(lib.ab_download_local(m.h, 1, m._ptr(vel), None, cap, C.byref(n)))
        return ids[: n.value], pos[: n.value], vel[: n.value]

    def n_owned(self):
        return int(self.model.lib.ab_nlocal(self.model.h))

    # -- host-buffer stepping (the end-to-end path of a sharded model) ---------------------------
    def host_state(self, slack=1.5):
        """Pinned host rows (global ids, pos, vel) of the agents this rank owns now."""
        import torch

        m = self.model
        ids, pos, vel = self.local_state()
        rows = int(len(ids) * slack) + 4096
        tdt = torch.float64 if m.space.dtype == np.float64 else torch.float32
        st = {"ids": torch.empty(rows, dtype=torch.int64, pin_memory=True).numpy(),
              "pos": torch.empty((rows, m.D), dtype=tdt, pin_memory=True).numpy(),
              "vel": torch.empty((rows, m.D), dtype=tdt, pin_memory=True).numpy(), "n": len(ids), "rows": rows}
        st["ids"][: len(ids)] = ids
        st["pos"][: len(ids)] = pos
        st["vel"][: len(ids)] = vel
        st["bytes_up"] = st["bytes_down"] = 0
        return st

    def host_step(self, st, nsteps=1):
        """upload the owned rows -> halo refresh from the neighbours (collective) -> step!(model, nsteps) -> download
        the rows owned afterwards (migration changes which)."""
        m, lib = self.model, self.model.lib
        n = st["n"]
        m._ck(lib.ab_upload_local(m.h, 0, m._ptr(st["pos"]), m._ptr(st["ids"]), n))
        m._ck(lib.ab_upload_local(m.h, 1, m._ptr(st["vel"]), m._ptr(st["ids"]), n))
        m._ck(lib.ab_comm_exchange(m.h))
        m.step(nsteps)
        k = C.c_int64(0)
        m._ck(lib.ab_download_local(m.h, 0, m._ptr(st["pos"]), m._ptr(st["ids"]), st["rows"], C.byref(k)))
        m._ck(lib.ab_download_local(m.h, 1, m._ptr(st["vel"]), None, st["rows"], C.byref(k)))
        row = (2 * m.D * st["pos"].itemsize + 8)
        st["bytes_up"], st["bytes_down"] = n * row, k.value * row
        st["n"] = k.value

    def gather(self):
        """Global id-ordered (pos, vel) on every rank (all_gather of the owned parts)."""
        ids, pos, vel = self.local_state()
        parts = [None] * self.world
        self.dist.all_gather_object(parts, (ids, pos, vel))
        return merge_parts(parts, self.n_global)


def merge_parts(parts, n_global):
    """Assemble per-rank (ids, pos, vel) into id-ordered arrays; every id must appear exactly once."""
    D = parts[0][1].shape[1]
    dt = parts[0][1].dtype
    pos = np.full((n_global, D), np.nan, dtype=dt)
    vel = np.full((n_global, D), np.nan, dtype=dt)
    seen = np.zeros(n_global, dtype=np.int64)
    for ids, p, v in parts:
        np.add.at(seen, ids - 1, 1)
        pos[ids - 1] = p
        vel[ids - 1] = v
    if not np.all(seen == 1):
        raise RuntimeError(f"agents lost or duplicated across ranks: {int((seen == 0).sum())} missing, "
                           f"{int((seen > 1).sum())} duplicated")
    return pos, vel


def init_slabs(model, dist, n_total, extent, state_fn, **kw):
    """bench.py helper: build the slabs and scatter the deterministic initial state."""
    s = Slabs(model, dist)
    pos, vel = state_fn(n_total, extent)
    s.scatter(pos, vel, **kw)
    return s


class LocalSlabs:
    """All slabs in ONE process (the single-session form a Julia host would use; also how the slab code is
    exercised on a single GPU): `nslabs` handles — on `devices[i]`, which may repeat — linked by
    ab_comm_link_local and stepped in lockstep by ab_step_flock_group."""

    def __init__(self, space_factory, rule, nslabs, devices=None):
        from . import _lib as L
        from .model import StandardABM

        self.L, self.rule, self.n = L, rule, int(nslabs)
        devices = list(devices) if devices is not None else [0] * self.n
        self.models = [StandardABM(space_factory(), model_step=rule, device=devices[r]) for r in range(self.n)]
        self.lib = self.models[0].lib
        self.slabs = []
        for r, m in enumerate(self.models):
            self.slabs.append(Slabs(m, dist=None, rank=r, world=self.n, init_nccl=False))
        self._hs = (C.c_void_p * self.n)(*[m.h for m in self.models])
        self.models[0]._ck(self.lib.ab_comm_link_local(self._hs, self.n))
        self.n_global = 0

    def scatter(self, pos, vel, **kw):
        self.n_global = len(pos)
        return [s.scatter(pos, vel, **kw) for s in self.slabs]

    def step(self, nsteps=1):
        r = self.rule
        fp = self.L.FlockParams(search=self.L.SEARCH[r.search], per_agent_field=-1, kernel=r.kernel, **r.p)
        self.models[0]._ck(self.lib.ab_step_flock_group(self._hs, self.n, C.byref(fp), int(nsteps)))

    def n_owned(self):
        return [s.n_owned() for s in self.slabs]

    def gather(self):
        return merge_parts([s.local_state() for s in self.slabs], self.n_global)

    def last_step_ms(self):
        return self.models[0].last_step_ms()

    def close(self):
        for m in self.models:
            m.close()
