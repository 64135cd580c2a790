"""Second, independent restatement of the two floating-point rules — brute force, O(N^2), numpy.

TEST INFRASTRUCTURE ONLY.  Written from the reference's example files and distance helpers directly
(NOT from oracle/abm_oracle.cpp): no cell list, no offset tables, no slot order.  Every bird looks at
every other bird.

  flock   examples/flock.jl:79-108 (agent_step!), src/spaces/utilities.jl:84-116 (euclidean_distance),
          :185-216 (get_direction), src/spaces/continuous.jl:221-238 (what `nearby_ids` returns),
          src/spaces/walk.jl:58-91 (walk! / normalize_position), synchronous update
          (every new velocity is computed from the OLD state, then everybody moves).
  pair    examples/celllistmap.jl:143-155 (calc_forces!), :176-191 (agent_step! integrator).

Neighbour sets are defined geometrically:
  :exact        j != i with euclidean_distance(pos_i, pos_j) <= visual_distance            (continuous.jl:237-238)
  :approximate  j != i whose CELL lies in the offset table of radius grid_r_i around bird i's cell:
                sum_d beta_d^2 <= grid_r_i^2 with beta = cell_j - cell_i taken to its nearest periodic image
                (grid_general.jl:138-151, continuous.jl:221-230) — cells come from pos2cell
                (continuous.jl:112-113), nothing else of the index is used.  Valid while the table is not
                clipped by the space (dims > 2 grid_r + 1 per periodic axis), which the callers guarantee.
Sums run in id order (numpy's pairwise summation), not in the cell-walk order of the reference, so
results agree with the oracle / the device to rounding (1e-13 relative in f64), while neighbour COUNTS
must agree exactly.
"""
import numpy as np


def _per(periodic, D):
    if isinstance(periodic, (bool, np.bool_)):
        periodic = (bool(periodic),) * D
    return np.asarray(periodic, dtype=bool)


def jl_mod(x, L):
    """Julia mod(x, y) for floats (result has the sign of y; may equal y after rounding)."""
    r = np.fmod(x, L)
    r = np.where((r != 0) & ((r < 0) != (L < 0)), r + L, r)
    return np.where(r == 0, np.copysign(0.0, L).astype(x.dtype), r).astype(x.dtype)


def normalize_position(x, extent, periodic):
    """walk.jl:77-91: mod on periodic axes, clamp(x, 0, prevfloat(L)) otherwise."""
    out = np.empty_like(x)
    for d in range(x.shape[1]):
        L = x.dtype.type(extent[d])
        if periodic[d]:
            out[:, d] = jl_mod(x[:, d], L)
        else:
            out[:, d] = np.clip(x[:, d], x.dtype.type(0), np.nextafter(L, x.dtype.type(-np.inf)))
    return out


def directions(pos, extent, periodic):
    """get_direction(pos_i, pos_j) for all (i, j): direct = to - from, inverse = direct - sign(direct) L,
    the one with the smaller |.| (ties: direct) on periodic axes (utilities.jl:185-216)."""
    T = pos.dtype.type
    direct = pos[None, :, :] - pos[:, None, :]
    out = direct.copy()
    for d in range(pos.shape[1]):
        if periodic[d]:
            inv = direct[:, :, d] - np.sign(direct[:, :, d]) * T(extent[d])
            out[:, :, d] = np.where(np.abs(direct[:, :, d]) <= np.abs(inv), direct[:, :, d], inv)
    return out


def distances(pos, extent, periodic):
    """euclidean_distance(pos_i, pos_j) for all (i, j) (utilities.jl:84-116)."""
    T = pos.dtype.type
    a = np.abs(pos[None, :, :] - pos[:, None, :])
    s = None
    for d in range(pos.shape[1]):
        m = np.minimum(a[:, :, d], T(extent[d]) - a[:, :, d]) if periodic[d] else a[:, :, d]
        s = m * m if s is None else s + m * m
    return np.sqrt(s)


def cells(pos, spacing):
    """pos2cell: max(1, ceil(Int, p / spacing)), 1-based (continuous.jl:112-113)."""
    T = pos.dtype.type
    return np.maximum(1, np.ceil(pos / T(spacing)).astype(np.int64))


def approximate_neighbours(pos, extent, spacing, periodic, r):
    """Boolean (i, j) matrix of nearby_ids_approx (continuous.jl:221-230), self excluded."""
    n, D = pos.shape
    c = cells(pos, spacing)
    dims = np.array([int(round(extent[d] / spacing)) for d in range(D)])
    # distance_from_cell_center: `cell .- 0.5` is Float64 whatever the eltype (continuous.jl:116-121)
    centre = spacing * (c.astype(np.float64) - 0.5)
    direct = np.abs(pos.astype(np.float64) - centre)
    delta2 = None
    for d in range(D):
        m = np.minimum(direct[:, d], extent[d] - direct[:, d]) if periodic[d] else direct[:, d]
        delta2 = m * m if delta2 is None else delta2 + m * m
    gr = np.ceil((r + np.sqrt(delta2)) / spacing).astype(np.int64)
    for d in range(D):
        assert (not periodic[d]) or dims[d] > 2 * gr.max() + 1, "offset table would be clipped: use a larger space"
    beta2 = np.zeros((n, n), dtype=np.int64)
    for d in range(D):
        b = c[None, :, d] - c[:, None, d]
        if periodic[d]:
            b = np.where(b > dims[d] // 2, b - dims[d], b)
            b = np.where(b < -(dims[d] // 2), b + dims[d], b)
        beta2 += b * b
    nb = beta2 <= (gr * gr)[:, None]
    np.fill_diagonal(nb, False)
    return nb


def flock_step(pos, vel, extent, spacing, periodic, p, search="exact"):
    """One synchronous flock step.  Returns (new_pos, new_vel, neighbour_counts)."""
    n, D = pos.shape
    T = pos.dtype.type
    per = _per(periodic, D)
    heading = directions(pos, extent, per)
    if search == "exact":
        nb = distances(pos, extent, per).astype(np.float64) <= float(p["visual_distance"])
        np.fill_diagonal(nb, False)
    else:
        nb = approximate_neighbours(pos, extent, spacing, per, float(p["visual_distance"]))
    s = None
    for d in range(D):
        sq = heading[:, :, d] * heading[:, :, d]
        s = sq if s is None else s + sq
    close = nb & (s < T(p["separation"]) * T(p["separation"]))          # flock.jl:93
    N = nb.sum(axis=1)
    cohere = np.where(nb[:, :, None], heading, T(0)).sum(axis=1)
    match = np.where(nb[:, :, None], vel[None, :, :], T(0)).sum(axis=1)
    separate = -np.where(close[:, :, None], heading, T(0)).sum(axis=1)
    cohere = cohere * T(p["cohere_factor"])
    separate = separate * T(p["separate_factor"])
    match = match * T(p["match_factor"])
    nv = vel + ((cohere + separate) + match) / np.maximum(N, 1).astype(pos.dtype)[:, None]
    nv = nv / np.sqrt((nv * nv).sum(axis=1))[:, None].astype(pos.dtype)
    npos = normalize_position(pos + T(p["speed"]) * nv, extent, per)       # move_agent!(bird, model, speed)
    return npos.astype(pos.dtype), nv.astype(pos.dtype), N


def pairforce_step(pos, vel, rad, k, mass, extent, periodic, dt, cutoff):
    """One step of the celllistmap example: forces over all unique pairs within `cutoff`, then the integrator."""
    n, D = pos.shape
    T = pos.dtype.type
    per = _per(periodic, D)
    dr = directions(pos, extent, per)                                      # y - x, minimum image
    d2 = None
    for d in range(D):
        sq = dr[:, :, d] * dr[:, :, d]
        d2 = sq if d2 is None else d2 + sq
    dist = np.sqrt(d2)
    rs = (rad[:, None] + rad[None, :]).astype(pos.dtype)
    act = (dist.astype(np.float64) <= float(cutoff)) & (dist <= rs)
    np.fill_diagonal(act, False)
    with np.errstate(divide="ignore", invalid="ignore"):
        pref = (T(2) * (k[:, None] * k[None, :]).astype(pos.dtype)) * (d2 - rs * rs)
        fij = np.where(act[:, :, None], pref[:, :, None] * (dr / dist[:, :, None]), T(0))
    f = fij.sum(axis=1)                                                    # forces[i] += fij over j (and -= from the other side)
    a = f / mass.astype(pos.dtype)[:, None]
    v = vel + a * T(dt)
    x = (pos + v * T(dt)) + (a / T(2)) * (T(dt) * T(dt))
    return normalize_position(x, extent, per).astype(pos.dtype), v.astype(pos.dtype)
