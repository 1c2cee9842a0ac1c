"""packages/lib/src/calculate.rs.  The std-dev that the sweep arg-maxes is computed on
the GPU in the reference's evaluation order (csrc/sweep.cu, sweep_finalize_kernel);
these two host helpers exist for callers that used the crate's functions directly on
small vectors.  They follow calculate.rs:2-23 literally (sequential f64)."""
import math


def get_arithmetic_mean(vec):              # calculate.rs:2-10
    s = float(vec[0])
    for x in vec[1:]:
        s = s + float(x)
    return s / float(len(vec))


def get_standard_deviation(vec):           # calculate.rs:13-23
    m = get_arithmetic_mean(vec)
    acc = (float(vec[0]) - m) * (float(vec[0]) - m)
    for x in vec[1:]:
        d = float(x) - m
        acc = acc + d * d
    return math.sqrt(acc / float(len(vec)))
