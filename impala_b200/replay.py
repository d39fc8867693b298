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
        self.draw_counter = None      # device-resident copy of `draws` (capturable_): for CUDA-graph replay

    def reset(self, seed: int) -> None:
        self.seed, self.draws = int(seed), 0
        if self.draw_counter is not None:
            self.draw_counter.zero_()

    def capturable_(self, device):
        """Keep the draw count in device memory so that ``ReplayBuffer.sample`` can be captured in a CUDA graph and replayed
        (a draw offset passed as a launch argument would be frozen into the graph).  Call outside capture."""
        if self.draw_counter is None:
            self.draw_counter = torch.tensor([self.draws], dtype=torch.int64, device=device)
        return self

    def sample(self, n: int, size: int, device, out=None) -> torch.Tensor:
        if size <= 0:
            raise RuntimeError("sample from an empty replay buffer")
        idx = ops.sample_uniform(self.seed, self.stream_id, self.draws, n, size, device, out)
        self.draws += n
        return idx


class PrioritizedSampler:
    """rlmeta ``PrioritizedSampler(priority_exponent)`` (reference ``agents/dqn/builder.py:83-90``): a device
    sum tree over the buffer slots; ``sample`` draws slots with probability ``p_i^alpha / sum`` and returns
    those probabilities (the learner turns them into importance weights, ``agents/dqn/learning.py:164-166``)."""

    prioritized = True

    def __init__(self, priority_exponent: float = 1.0, stream_id: int = 0):
        self.alpha, self.seed, self.stream_id, self.draws = float(priority_exponent), 0, int(stream_id), 0
        self.tree = None

    def reset(self, seed: int) -> None:
        self.seed, self.draws = int(seed), 0

    def bind(self, capacity: int, device) -> None:
        leaves = 1
        while leaves < capacity:
            leaves *= 2
        self.tree = torch.zeros(2 * leaves, dtype=torch.float64, device=device)

    def update(self, slots: torch.Tensor, priorities: torch.Tensor) -> None:
        value = priorities.to(self.tree.device, torch.float64).reshape(-1)
        if self.alpha != 1.0:
            value = value.pow(self.alpha)
        ops.sumtree_update(self.tree, slots.to(self.tree.device).reshape(-1), value)

    def sample(self, n: int, size: int, device, out=None):
        if size <= 0:
            raise RuntimeError("sample from an empty replay buffer")
        idx, prob = ops.sumtree_sample(self.tree, size, self.seed, self.stream_id, self.draws, n)
        self.draws += n
        return idx, prob


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
    def _prioritized(self) -> bool:
        return getattr(self.sampler, "prioritized", False)

    def _set_priorities(self, keys: range, priorities) -> None:
        if not self._prioritized():
            return
        if self.sampler.tree is None:
            self.sampler.bind(self.storage.capacity, self.storage.device)
        n = len(keys)
        if priorities is None:                       # rlmeta: new items enter with priority 1
            priorities = torch.ones(n, dtype=torch.float64, device=self.storage.device)
        priorities = torch.as_tensor(priorities).reshape(-1)[-n:]           # rows beyond capacity were dropped
        slots = torch.arange(keys.start, keys.stop, device=self.storage.device) % self.storage.capacity
        self.sampler.update(slots, priorities)

    def append(self, item: Sequence[torch.Tensor], priority=None):
        """One item = one list of per-field tensors (IMPALA: a whole ``[T,...]`` rollout)."""
        keys = self.storage.extend_stacked([f.unsqueeze(0) for f in item])
        self._set_priorities(keys, None if priority is None else torch.as_tensor([priority]))
        return keys

    def extend(self, items, priorities=None):
        """A list of items (each a list of per-field tensors), or a tuple of stacked fields; ``priorities`` as at
        reference ``agents/dqn/learning.py:54`` (prioritized sampler only)."""
        if isinstance(items, tuple):
            keys = self.storage.extend_stacked(list(items))
        elif len(items) == 0:
            return range(0)
        else:
            keys = self.storage.extend_stacked([torch.stack(f) for f in zip(*items)])
        self._set_priorities(keys, priorities)
        return keys

    def update(self, keys: torch.Tensor, priorities: torch.Tensor) -> None:
        """``replay_buffer.update(keys, priorities)`` (reference ``agents/dqn/learning.py:184``): keys of rows
        that have been overwritten since they were sampled are ignored."""
        if not self._prioritized():
            return
        st = self.storage
        keys = keys.to(st.device).reshape(-1)
        live = keys >= max(st.count - st.capacity, 0)
        slots = torch.where(live, keys % st.capacity, torch.full_like(keys, -1))   # -1: skipped by the kernel
        self.sampler.update(slots, priorities)

    def extend_rollouts(self, s, a, r, done, pi_ref, gamma: float = 0.99):
        """Batched ``ImpalaActor._make_replay`` + ``async_append`` for ``n`` stacked rollouts ``[n,T,...]``
        (reference ``agents/impala/learning.py:71-80``): stores ``[s, a, r, logical_not(done)*gamma, pi_ref]``."""
        return self.storage.extend_rollouts([s, a, r, done, pi_ref], done_field=3, gamma=gamma)

    # -- learner side (reference agents/impala/learning.py:117,121) -------------------------------
    def sample(self, n: int):
        """``(keys, batch, probabilities)``; ``batch`` = list of stacked ``[n, ...]`` device tensors."""
        st = self.storage
        size = st.size
        if self._prioritized():
            slots, probs = self.sampler.sample(n, size, st.device)
            batch = st.gather(slots)
        elif type(self.sampler) is UniformSampler and st.fields is not None and len(st.fields) <= 8:
            # uniform sampler: the draw happens inside the gather launch (K4 + K5 in one kernel, same Philox stream)
            probs = None
            ctr = self.sampler.draw_counter
            slots, batch = ops.sample_gather_fields(st.fields, self.sampler.seed, self.sampler.stream_id,
                                                    0 if ctr is not None else self.sampler.draws, n, size, draw_counter=ctr)
            if ctr is not None:
                ops.counter_add(ctr, n)       # in stream order, after the draw: the next (replayed) sample continues the stream
            self.sampler.draws += n
        else:
            slots, probs = self.sampler.sample(n, size, st.device), None
            batch = st.gather(slots)
        if not self.with_keys:
            return slots, batch, None
        oldest = max(st.count - st.capacity, 0)
        keys = oldest + torch.remainder(slots - oldest, st.capacity)
        if probs is None:
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
