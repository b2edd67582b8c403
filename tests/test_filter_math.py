"""CPU check of the MATHEMATICS of the tensor-core distance filter (csrc/psim_allpairs_tc.cuh: filter_record_tc,
hsplit2 / hsplit3, the K = 16 FP16 operand layout with the cut folded in, FP16 accumulator, sign test), restated
in numpy.

What it proves: for every ordered pair with true r < r_cut the accumulator entry stays negative (and above FP16's
smallest subnormal in magnitude) even under a worst-case model of the tensor core's arithmetic — the products of
the 11-bit FP16 pieces are exact (22 bits), and the 16-term accumulation is charged 16 roundings of 2^-23 relative
to the sum of the magnitudes (the hardware measured far better: tools/microbench/tc_f16.cu) — on the reference's
shipped inputs, in the corners of large boxes and in a tiny box (large scale factor).  What it does not prove:
that the CUDA code computes this (the GPU tests do that through psim_counters.filter_pairs ==
psim_count_pairs_in_range).  No GPU, no oracle.
"""
import os

import numpy as np
import pytest
from scipy.spatial import cKDTree

from conftest import GOLDEN

U32 = np.float32(5.9604645e-8)          # 2^-24: unit roundoff of the float coordinates, as in filter_record_tc
R_CUT = 2.5


def _lattice(n_side, origin, jitter, seed):
    rng = np.random.default_rng(seed)
    g = np.stack(np.meshgrid(*[np.arange(n_side)] * 3, indexing="ij"), -1).reshape(-1, 3) * 2.0 ** (1.0 / 6.0)
    return g + np.asarray(origin) + rng.uniform(-jitter, jitter, g.shape)


BIG = 60000.0


def scale16(area):
    """SimT::init: the largest power of two S with (1.4 x half diagonal x S)^2 <= 60000."""
    lim = np.sqrt(BIG) / (1.4 * 0.5 * np.sqrt(np.sum(np.square(area))))
    return np.float32(2.0 ** np.floor(np.log2(lim)))


def hsplit2(x):
    a = np.asarray(x, np.float32).astype(np.float16)
    return a, (np.asarray(x, np.float32) - a.astype(np.float32)).astype(np.float16)


def hsplit3(x):
    x = np.asarray(x, np.float32)
    a = x.astype(np.float16)
    r = x - a.astype(np.float32)
    b = r.astype(np.float16)
    return a, b, (r - b.astype(np.float32)).astype(np.float16)


def worst_case_margin16(pos, area):
    """min over in-range ordered pairs of -(acc_ij + accumulation error bound): the accumulator must stay
    NEGATIVE (and representable in FP16: magnitude above 2^-24) for the pair to reach the exact path.  In units
    of the scaled r_cut^2."""
    S = scale16(area)
    S2 = S * S
    eps = np.float64(np.finfo(np.float64).eps)
    cut2_pad = np.float32(R_CUT * R_CUT * (1.0 + 32.0 * eps))
    cut2s = cut2_pad * np.float32(1.0 + 1.0e-6) * S2
    f = (pos - 0.5 * np.asarray(area)).astype(np.float32)                       # centred, in the working precision first
    n2 = (f[:, 0] * f[:, 0] + f[:, 1] * f[:, 1] + f[:, 2] * f[:, 2]).astype(np.float32)
    ys = f * S
    n2s = n2 * S2
    w = (n2s * np.float32(1.0 - 2.0 ** -16) - S2 * np.float32(16.0) * U32 * np.float32(R_CUT) * (np.sqrt(n2) + np.float32(1.0))
         - np.float32(2.0 ** -13)).astype(np.float32)
    assert (n2s < 0.9 * BIG).all()
    pairs = cKDTree(pos).query_pairs(R_CUT, output_type="ndarray")
    d = pos[pairs[:, 0]] - pos[pairs[:, 1]]
    pairs = pairs[np.sqrt((d * d).sum(-1)) < R_CUT]
    pairs = np.concatenate([pairs, pairs[:, ::-1]])
    i, j = pairs[:, 0], pairs[:, 1]
    terms = []
    for k in range(3):
        a1, a2 = hsplit2(ys[:, k])
        b1, b2 = hsplit2(np.float32(-2.0) * ys[:, k])
        terms += [a1[i].astype(np.float64) * b1[j].astype(np.float64), a1[i].astype(np.float64) * b2[j].astype(np.float64),
                  a2[i].astype(np.float64) * b1[j].astype(np.float64)]
    w1, w2, w3 = hsplit3(w)
    c1, c2, c3 = hsplit3(-(cut2s - w))
    terms += [w1[j].astype(np.float64), w2[j].astype(np.float64), w3[j].astype(np.float64),
              c1[i].astype(np.float64), c2[i].astype(np.float64), c3[i].astype(np.float64)]
    terms = np.stack(terms)
    assert np.isfinite(terms).all()
    acc = terms.sum(0)
    acc_err = 16.0 * 2.0 ** -23 * np.abs(terms).sum(0)
    worst = -(acc + acc_err)                     # must be positive, and survive FP16's smallest subnormal
    return (worst.min() - 2.0 ** -24) / (R_CUT * R_CUT * float(S2)), len(i), float(S)


@pytest.mark.parametrize("case", ["shipped_2000", "shipped_39440", "corner_of_a_198_box", "corner_of_a_1000_box", "dense_jitter",
                                  "tiny_box"])
def test_fp16_tensor_core_filter_is_conservative_under_worst_case_rounding(case):
    area = (99.0, 99.0, 99.0)
    if case == "shipped_2000":
        pos = np.load(os.path.join(GOLDEN, "shipped_test_particles_golden.npz"))["x0"]
    elif case == "shipped_39440":
        pos = np.round(np.load(os.path.join(GOLDEN, "shipped_more_test_particles_golden.npz"))["x0"].astype(np.float64), 4)
    elif case == "corner_of_a_198_box":
        pos, area = _lattice(16, (175.0, 175.0, 80.0), 0.05, 1), (198.0, 198.0, 99.0)
    elif case == "corner_of_a_1000_box":
        pos, area = _lattice(14, (980.0, 975.0, 980.0), 0.05, 2), (1002.0, 1002.0, 1002.0)
    elif case == "tiny_box":                     # large S: the pieces of small coordinates must not underflow harmfully
        pos, area = _lattice(8, (0.2, 0.2, 0.2), 0.05, 4), (12.0, 12.0, 12.0)
    else:
        pos = _lattice(14, (40.0, 40.0, 40.0), 0.45, 3)
    margin, n_pairs, S = worst_case_margin16(pos, area)
    assert n_pairs > 4000
    assert margin > 0.0, "an in-range pair could be dropped by the FP16 filter (S = %g)" % S
