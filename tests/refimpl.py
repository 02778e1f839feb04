"""Independent numpy model of the generalized allreduce (written from the algorithm definition in
SURVEY.md 2.7, not from the C++): used as the oracle for the CPU world and the GPU kernels."""
import numpy as np


def block_size(count, n, esize, vectorized=True):
    b = max(1, -(-count // n))
    vec = 16 // esize if vectorized else 1
    return -(-b // vec) * vec


def staged_allreduce(inputs, widths, op=np.add, block=None, round_fn=None, final_scale=None):
    """inputs: list of N 1-D arrays (same dtype/len). Summation order inside a stage: own value first,
    then the group peers in ascending rank order. round_fn rounds stage results to the wire dtype."""
    n = len(inputs)
    count = inputs[0].size
    p = 1
    for w in widths:
        p *= w
    assert p == n
    B = block if block is not None else max(1, -(-count // n))
    pad = n * B
    acc_dtype = np.float32 if inputs[0].dtype in (np.float16,) else inputs[0].dtype
    W = [np.zeros(pad, dtype=acc_dtype) for _ in range(n)]
    for r in range(n):
        W[r][:count] = inputs[r].astype(acc_dtype)
    gaps = [1]
    for w in widths:
        gaps.append(gaps[-1] * w)

    def blk(b):
        return slice(b * B, (b + 1) * B)

    for k, w in enumerate(widths):  # reduce-scatter
        g, gn = gaps[k], gaps[k + 1]
        new = [x.copy() for x in W]
        for r in range(n):
            base = r - ((r // g) % w) * g
            peers = [base + j * g for j in range(w)]
            for b in range(r % gn, n, gn):
                acc = W[r][blk(b)].copy()
                for pr in peers:
                    if pr != r:
                        acc = op(acc, W[pr][blk(b)])
                if final_scale is not None and k == len(widths) - 1:
                    acc = acc * acc.dtype.type(final_scale)  # scaled once, when the block becomes final
                if round_fn is not None:
                    acc = round_fn(acc)
                new[r][blk(b)] = acc
        W = new
    for k in range(len(widths) - 1, -1, -1):  # all-gather
        w, g, gn = widths[k], gaps[k], gaps[k + 1]
        new = [x.copy() for x in W]
        for r in range(n):
            base = r - ((r // g) % w) * g
            for j in range(w):
                pr = base + j * g
                for b in range(pr % gn, n, gn):
                    new[r][blk(b)] = W[pr][blk(b)]
        W = new
    return [x[:count] for x in W]


def to_bf16_bits(x):
    """float32 array -> uint16 bf16 bits, round to nearest even."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
    return r


def from_bf16_bits(b):
    return (b.astype(np.uint32) << 16).view(np.float32)


def round_bf16(x):
    return from_bf16_bits(to_bf16_bits(x))
