"""Specification + CPU restatement of the replay sample/gather path (TEST ORACLE).

Replaces rlmeta ``ReplayBuffer(CircularBuffer(capacity), UniformSampler())``
(reference ``agents/impala/builder.py:30-36``, consumer
``agents/impala/learning.py:121``).  rlmeta's C++ RNG stream is unavailable
(PARITY UNPINNED), so the stream is specified here and pinned by the Random123
Philox4x32-10 known-answer vectors (``tests/test_oracle_replay.py``):

* draw ``g`` (64-bit global draw number = draws made before this call + i) of
  stream ``(seed, stream_id)``:
  ``ctr = (lo32(g), hi32(g), stream_id, 0)``, ``key = (lo32(seed), hi32(seed))``;
  ``x64 = out[0] | out[1] << 32``; ``index = mulhi64(x64, size)``
  (with replacement; bias < size / 2**64).
* ring buffer: key = append count (monotone), slot = key mod capacity,
  probabilities = 1/size.
"""
import numpy as np

M0, M1 = 0xD2511F53, 0xCD9E8D57
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK32 = 0xFFFFFFFF


def philox4x32_10(ctr, key):
    """Scalar Philox4x32-10 (Salmon et al. 2011, Random123).  ``ctr``: 4 uint32, ``key``: 2."""
    c0, c1, c2, c3 = [int(x) & MASK32 for x in ctr]
    k0, k1 = [int(x) & MASK32 for x in key]
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        c0, c1, c2, c3 = ((p1 >> 32) ^ c1 ^ k0) & MASK32, p1 & MASK32, \
                         ((p0 >> 32) ^ c3 ^ k1) & MASK32, p0 & MASK32
        k0 = (k0 + W0) & MASK32
        k1 = (k1 + W1) & MASK32
    return c0, c1, c2, c3


def philox4x32_10_vec(c0, c1, c2, c3, k0, k1):
    """Vectorised over numpy uint64 arrays holding 32-bit values."""
    c0, c1, c2, c3 = [np.asarray(x, dtype=np.uint64) for x in (c0, c1, c2, c3)]
    k0 = np.uint64(k0)
    k1 = np.uint64(k1)
    m32 = np.uint64(MASK32)
    s32 = np.uint64(32)
    for _ in range(10):
        p0 = np.uint64(M0) * c0
        p1 = np.uint64(M1) * c2
        c0, c1, c2, c3 = ((p1 >> s32) ^ c1 ^ k0) & m32, p1 & m32, \
                         ((p0 >> s32) ^ c3 ^ k1) & m32, p0 & m32
        k0 = (k0 + np.uint64(W0)) & m32
        k1 = (k1 + np.uint64(W1)) & m32
    return c0, c1, c2, c3


def _mulhi64(x, n):
    """floor(x * n / 2**64) for uint64 arrays x and python int n < 2**63."""
    x = np.asarray(x, dtype=np.uint64)
    n = int(n)
    m32 = np.uint64(MASK32)
    s32 = np.uint64(32)
    xl, xh = x & m32, x >> s32
    nl, nh = np.uint64(n & MASK32), np.uint64(n >> 32)
    ll = xl * nl
    lh = xl * nh
    hl = xh * nl
    hh = xh * nh
    mid = (ll >> s32) + (lh & m32) + (hl & m32)
    return hh + (lh >> s32) + (hl >> s32) + (mid >> s32)


def sample_uniform(seed, stream_id, draw_offset, n, size):
    """``n`` indices in ``[0, size)`` for draws ``draw_offset .. draw_offset+n-1``. int64 array."""
    if n == 0:
        return np.zeros(0, dtype=np.int64)
    if size <= 0:
        raise RuntimeError("sample from empty buffer")
    g = np.arange(n, dtype=np.uint64) + np.uint64(draw_offset)
    o0, o1, _, _ = philox4x32_10_vec(g & np.uint64(MASK32), g >> np.uint64(32),
                                     np.full(n, stream_id & MASK32, dtype=np.uint64),
                                     np.zeros(n, dtype=np.uint64),
                                     seed & MASK32, (seed >> 32) & MASK32)
    x64 = o0 | (o1 << np.uint64(32))
    return _mulhi64(x64, size).astype(np.int64)


class UniformSampler:
    """rlmeta ``UniformSampler`` stand-in: ``reset(seed)`` (``agents/impala/builder.py:31-32``)."""

    def __init__(self, stream_id=0):
        self.seed, self.stream_id, self.draws = 0, stream_id, 0

    def reset(self, seed):
        self.seed, self.draws = int(seed), 0

    def sample(self, n, size):
        idx = sample_uniform(self.seed, self.stream_id, self.draws, n, size)
        self.draws += n
        return idx


class CircularBuffer:
    """Fixed-capacity ring of per-field numpy rows; slot = key mod capacity."""

    def __init__(self, capacity):
        self.capacity, self.count, self.fields = int(capacity), 0, None

    @property
    def size(self):
        return min(self.count, self.capacity)

    def append(self, item):
        return self.extend([[np.asarray(f)[None] for f in item]], _stacked=True)

    def extend(self, items, _stacked=False):
        """``items``: list of per-field arrays ``[n, ...]`` (already stacked along dim 0)."""
        fields = items[0] if _stacked else items
        n = fields[0].shape[0]
        if self.fields is None:
            self.fields = [np.zeros((self.capacity,) + f.shape[1:], dtype=f.dtype) for f in fields]
        keys = np.arange(self.count, self.count + n, dtype=np.int64)
        for k, j in zip(keys, range(n)):       # in order: later rows overwrite earlier ones
            for dst, f in zip(self.fields, fields):
                dst[k % self.capacity] = f[j]
        self.count += n
        return keys

    def gather(self, slots):
        return [f[slots] for f in self.fields]


class ReplayBuffer:
    """``sample(n) -> (keys, batch, probabilities)`` (``agents/impala/learning.py:121``)."""

    def __init__(self, storage, sampler):
        self.storage, self.sampler = storage, sampler

    def append(self, item):
        return self.storage.append(item)

    def extend(self, fields):
        return self.storage.extend(fields)

    def sample(self, n):
        size = self.storage.size
        slots = self.sampler.sample(n, size)
        oldest = max(self.storage.count - self.storage.capacity, 0)
        # key of the row living in slot s: the unique k in [oldest, count) with k % cap == s
        keys = oldest + ((slots - oldest) % self.storage.capacity)
        probs = np.full(n, 1.0 / size, dtype=np.float32) if n else np.zeros(0, np.float32)
        return keys, self.storage.gather(slots), probs


# --------------------------------------------------------------------------------------------
# Prioritized sampling: specification of the device sum tree.
# rlmeta PrioritizedSampler (reference agents/dqn/builder.py:83-90) is unobtainable -> PARITY UNPINNED;
# what is pinned is this specification, reproduced bit for bit by impala_b200/csrc/sumtree.cu.
# --------------------------------------------------------------------------------------------
class SumTree:
    """Complete binary tree of float64 sums over ``leaves`` (power of two) slots; node i = node 2i + node 2i+1."""

    def __init__(self, capacity):
        self.leaves = 1
        while self.leaves < capacity:
            self.leaves *= 2
        self.tree = np.zeros(2 * self.leaves, dtype=np.float64)

    def update(self, idx, value):
        idx = np.asarray(idx, dtype=np.int64)
        self.tree[self.leaves + idx] = np.asarray(value, dtype=np.float64)
        nodes = np.unique((self.leaves + idx) >> 1)
        while nodes.size and nodes[0] >= 1:
            self.tree[nodes] = self.tree[2 * nodes] + self.tree[2 * nodes + 1]
            if nodes[0] == 1:
                break
            nodes = np.unique(nodes >> 1)

    def sample(self, seed, stream_id, draw_offset, n, size):
        g = np.arange(n, dtype=np.uint64) + np.uint64(draw_offset)
        o0, o1, _, _ = philox4x32_10_vec(g & np.uint64(MASK32), g >> np.uint64(32),
                                         np.full(n, stream_id & MASK32, dtype=np.uint64), np.zeros(n, dtype=np.uint64),
                                         seed & MASK32, (seed >> 32) & MASK32)
        x64 = o0 | (o1 << np.uint64(32))
        total = self.tree[1]
        mass = ((x64 >> np.uint64(11)).astype(np.float64) * 2.0 ** -53) * total
        out = np.zeros(n, dtype=np.int64)
        for i in range(n):
            node, m = 1, mass[i]
            while node < self.leaves:
                left = self.tree[2 * node]
                if m < left:
                    node = 2 * node
                else:
                    m = m - left
                    node = 2 * node + 1
            out[i] = min(node - self.leaves, size - 1)
        prob = (self.tree[self.leaves + out] / total).astype(np.float32)
        return out, prob
