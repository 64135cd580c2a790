"""Known answers computed by hand from the reference's source, for the two rows of the path that the reference's own
test-suite holds no vector for.  Both the CPU restatement (tests/test_oracle_golden.py) and the
device (tests/test_field_gpu.py, tests/test_pairforce_gpu.py) are checked against these numbers.

Pair force — examples/celllistmap.jl:143-155 (calc_forces!) and :176-191 (agent_step!), periodic (100, 100) space:
  for a pair (i, j) with d = |dr| <= r_i + r_j, dr = x_j - x_i (minimum image):
      fij = 2 (k_i k_j) (d^2 - (r_i + r_j)^2) dr / d;   forces[i] += fij;   forces[j] -= fij
  then per particle  a = f / mass;  v = vel + a dt;  x = pos + v dt + (a / 2) dt^2   (x uses the NEW v, as written).

  particles            pos        vel      r    k    mass
    1               (10, 10)    (1, 0)     4    1     2
    2               (13, 14)    (0, -1)    4    1     4
    3               (98, 10)    (0, 0)     9   0.5    5
  pair (1, 2): dr = (3, 4), d = 5 <= 8:    fij = 2 * 1 * (25 - 64) * (0.6, 0.8) = (-46.8, -62.4)
  pair (1, 3): dr = (-12, 0) across the periodic edge, d = 12 <= 13:
                                           fij = 2 * 0.5 * (144 - 169) * (-1, 0) = (25, 0)
  pair (2, 3): dr = (-15, -4), d = 15.52 > 13: no force.
  two-particle system (1, 2 only), dt = 0.1:
      f1 = (-46.8, -62.4): a1 = (-23.4, -31.2), v1 = (-1.34, -3.12), x1 = (10 - 0.134 - 0.117, 10 - 0.312 - 0.156)
      f2 = (46.8, 62.4):   a2 = (11.7, 15.6),   v2 = (1.17, 0.56),   x2 = (13 + 0.117 + 0.0585, 14 + 0.056 + 0.078)
  three-particle system, dt = 0.1:
      f1 = (-21.8, -62.4): a1 = (-10.9, -31.2), v1 = (-0.09, -3.12), x1 = (10 - 0.009 - 0.0545, 9.532)
      f2 as above;  f3 = (-25, 0): a3 = (-5, 0), v3 = (-0.5, 0), x3 = (98 - 0.05 - 0.025, 10)

Spatial index — src/spaces/continuous.jl:566-573: eps = spacesize[1:D] ./ size(property), idx = floor.(Int, pos[1:D] ./ eps) .+ 1;
property[idx] throws BoundsError when pos == extent (idx = size + 1).
"""
import numpy as np

PF_EXTENT, PF_SPACING, PF_CUTOFF, PF_DT = (100.0, 100.0), 20.0, 20.0, 0.1
PF_POS = np.array([[10.0, 10.0], [13.0, 14.0], [98.0, 10.0]])
PF_VEL = np.array([[1.0, 0.0], [0.0, -1.0], [0.0, 0.0]])
PF_R = np.array([4.0, 4.0, 9.0])
PF_K = np.array([1.0, 1.0, 0.5])
PF_MASS = np.array([2.0, 4.0, 5.0])
PF2_POS1 = np.array([[9.749, 9.532], [13.1755, 14.134]])
PF2_VEL1 = np.array([[-1.34, -3.12], [1.17, 0.56]])
PF3_POS1 = np.array([[9.9365, 9.532], [13.1755, 14.134], [97.925, 10.0]])
PF3_VEL1 = np.array([[-0.09, -3.12], [1.17, 0.56], [-0.5, 0.0]])
PF_TOL = 1e-12          # the hand values are decimal literals; the arithmetic (dr / d = 0.6, 0.8) is not exact in binary

# space (10, 6); property of size (5, 3): eps = (2, 2);  property of size (4,): eps = 2.5 on the first coordinate
SI_EXTENT = (10.0, 6.0)
SI_POS = np.array([[0.0, 0.0], [2.0, 1.999], [3.999, 4.0], [9.99, 5.99], [2.5, 3.0], [7.4999, 0.1], [7.5, 5.0]])
SI_IDX_5x3 = np.array([[1, 1], [2, 1], [2, 3], [5, 3], [2, 2], [4, 1], [4, 3]])
SI_IDX_4 = np.array([[1], [1], [2], [4], [2], [3], [4]])
SI_POS_AT_EXTENT = np.array([[10.0, 3.0]])          # idx = (6, 2): outside a (5, 3) property — BoundsError in the reference
