"""Device-resident replay: drop-in for the rlmeta objects the reference builds at
``agents/impala/builder.py:30-36`` / ``agents/ppo/builder.py:30-36``::

    sampler = UniformSampler(); sampler.reset(seed)
    rb = ReplayBuffer(CircularBuffer(capacity), sampler)
    rb.append(item) / rb.extend(items);  keys, batch, probs = rb.sample(n)

Storage is one HBM tensor per field (``[capacity, *field_shape]``, allocated on the first
append), indices are drawn on the device (Philox4x32-10, bit-exact against
``oracle/replay_ref.py``) and ``sample`` gathers rows HBM->HBM straight into stacked
``[n, ...]`` batch tensors -- i.e. what the reference gets only after its RPC, 5*B
host-to-device copies and ``collate_nested(torch.stack)`` (``agents/impala/learning.py:121-123,142``).
"""
from typing import List, Optional, Sequence

import torch

from . import ops


class UniformSampler:
    """rlmeta ``UniformSampler``: ``reset(seed)``; with replacement; one Philox stream per
    ``(seed, stream_id)`` (``stream_id`` = rank for rank-local buffers)."""

    def __init__(self, stream_id: int = 0):
        self.seed, self.stream_id, self.draws = 0, int(stream_id), 0

    def reset(self, seed: int) -> None:
        self.seed, self.draws = int(seed), 0

    def sample(self, n: int, size: int, device, out=None) -> torch.Tensor:
        if size <= 0:
            raise RuntimeError("sample from an empty replay buffer")
        idx = ops.sample_uniform(self.seed, self.stream_id, self.draws, n, size, device, out)
        self.draws += n
        return idx


class CircularBuffer:
    """rlmeta ``CircularBuffer(capacity, collate_fn=None)``: key = append count, slot = key % capacity."""

    def __init__(self, capacity: int, collate_fn=None, device="cuda"):
        self.capacity, self.count = int(capacity), 0
        self.collate_fn = collate_fn  # accepted for signature compatibility; rows are always stacked
        self.device = torch.device(device)
        self.fields: Optional[List[torch.Tensor]] = None

    @property
    def size(self) -> int:
        return min(self.count, self.capacity)

    def _alloc(self, rows: Sequence[torch.Tensor]) -> None:
        self.fields = [torch.empty((self.capacity,) + tuple(r.shape[1:]), dtype=r.dtype, device=self.device)
                       for r in rows]

    def extend_stacked(self, rows: Sequence[torch.Tensor]) -> range:
        """``rows``: one ``[n, *field_shape]`` tensor per field (host or device)."""
        n = rows[0].shape[0]
        if n > self.capacity:  # only the last `capacity` rows can survive
            rows = [r[n - self.capacity:] for r in rows]
            self.count += n - self.capacity
            n = self.capacity
        rows = [r.to(self.device, non_blocking=True) for r in rows]
        if self.fields is None:
            self._alloc(rows)
        if len(rows) <= 8:
            ops.ring_append_fields(self.fields, self.count, rows)           # one launch for all fields
        else:
            for dst, r in zip(self.fields, rows):
                ops.ring_append(dst, self.count, r)
        keys = range(self.count, self.count + n)
        self.count += n
        return keys

    def gather(self, slots: torch.Tensor) -> List[torch.Tensor]:
        if len(self.fields) <= 8:
            return ops.gather_fields(self.fields, slots)                    # one launch for all fields
        return [ops.gather_rows(f, slots) for f in self.fields]

    def extend_rollouts(self, rows: Sequence[torch.Tensor], done_field: int, gamma: float) -> range:
        """Device-side ``_make_replay`` + append (reference ``agents/impala/learning.py:71-80``): ``rows`` are the
        stacked actor outputs ``[n, T, ...]`` with bool ``done`` flags at ``done_field``; the stored field is
        ``discount_t = logical_not(done) * gamma`` (float32), computed inside the one append launch."""
        n = rows[0].shape[0]
        if n > self.capacity:
            rows = [r[n - self.capacity:] for r in rows]
            self.count += n - self.capacity
            n = self.capacity
        rows = [r.to(self.device, non_blocking=True) for r in rows]
        if self.fields is None:
            self.fields = [torch.empty((self.capacity,) + tuple(r.shape[1:]),
                                       dtype=torch.float32 if i == done_field else r.dtype, device=self.device)
                           for i, r in enumerate(rows)]
        ops.ring_append_fields(self.fields, self.count, rows, done_field=done_field, gamma=gamma)
        keys = range(self.count, self.count + n)
        self.count += n
        return keys


class ReplayBuffer:
    """rlmeta ``ReplayBuffer(storage, sampler)`` surface used by the reference learners."""

    def __init__(self, storage: CircularBuffer, sampler: UniformSampler, with_keys: bool = True):
        self.storage, self.sampler, self.with_keys = storage, sampler, with_keys

    # -- writer side (reference agents/impala/learning.py:75, agents/ppo/learning.py:63) ----------
    def append(self, item: Sequence[torch.Tensor]):
        """One item = one list of per-field tensors (IMPALA: a whole ``[T,...]`` rollout)."""
        return self.storage.extend_stacked([f.unsqueeze(0) for f in item])

    def extend(self, items):
        """A list of items (each a list of per-field tensors), or a tuple of stacked fields."""
        if isinstance(items, tuple):
            return self.storage.extend_stacked(list(items))
        if len(items) == 0:
            return range(0)
        return self.storage.extend_stacked([torch.stack(f) for f in zip(*items)])

    def extend_rollouts(self, s, a, r, done, pi_ref, gamma: float = 0.99):
        """Batched ``ImpalaActor._make_replay`` + ``async_append`` for ``n`` stacked rollouts ``[n,T,...]``
        (reference ``agents/impala/learning.py:71-80``): stores ``[s, a, r, logical_not(done)*gamma, pi_ref]``."""
        return self.storage.extend_rollouts([s, a, r, done, pi_ref], done_field=3, gamma=gamma)

    # -- learner side (reference agents/impala/learning.py:117,121) -------------------------------
    def sample(self, n: int):
        """``(keys, batch, probabilities)``; ``batch`` = list of stacked ``[n, ...]`` device tensors."""
        st = self.storage
        size = st.size
        slots = self.sampler.sample(n, size, st.device)
        batch = st.gather(slots)
        if not self.with_keys:
            return slots, batch, None
        oldest = max(st.count - st.capacity, 0)
        keys = oldest + torch.remainder(slots - oldest, st.capacity)
        probs = torch.full((n,), 1.0 / size, dtype=torch.float32, device=st.device)
        return keys, batch, probs

    def warm_up(self, learning_starts: Optional[int] = None) -> None:
        """The reference blocks here until actors have filled ``learning_starts`` items
        (``agents/impala/learning.py:117``); with a device-local buffer that is a precondition."""
        if learning_starts is not None and self.storage.size < min(learning_starts, self.storage.capacity):
            raise RuntimeError("replay buffer holds %d items, learning_starts=%d"
                               % (self.storage.size, learning_starts))

    def info(self):
        return self.storage.capacity, self.storage.size

    @property
    def size(self) -> int:
        return self.storage.size

    @property
    def capacity(self) -> int:
        return self.storage.capacity

    def __len__(self) -> int:
        return self.storage.size
