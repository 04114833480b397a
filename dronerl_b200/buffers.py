"""Device-resident replay ring buffer -- the consumer right after the env hot path.

Mirrors /root/reference/jax_impl/buffers.py:17-93 (``init / add / add_many / sample / can_sample``,
ring indices ``(current_idx + arange(k)) % buffer_size``) on torch tensors.  ``current_idx`` and
``current_size`` live on the device as 0-d int64 tensors and are updated with tensor ops only, so
``add_many`` and ``sample`` can sit inside a captured CUDA graph next to the env kernels.  Plain
PyTorch on purpose: this is plumbing around the path, not the path.
"""
from dataclasses import dataclass
from typing import Dict

import torch


@dataclass
class BufferState:
    experiences: Dict[str, torch.Tensor]
    current_idx: torch.Tensor      # 0-d int64
    current_size: torch.Tensor     # 0-d int64

    def replace(self, **kw):
        d = dict(experiences=self.experiences, current_idx=self.current_idx, current_size=self.current_size)
        d.update(kw)
        return BufferState(**d)


class ReplayBuffer:
    def __init__(self, buffer_size: int = 10_000, sample_batch_size: int = 64):
        self.buffer_size = int(buffer_size)
        self.sample_batch_size = int(sample_batch_size)

    def init(self, experience: Dict[str, torch.Tensor]) -> BufferState:
        """buffers.py:26-38: one example experience fixes the per-field shape and dtype."""
        exps = {k: torch.zeros((self.buffer_size, *torch.as_tensor(v).shape), dtype=torch.as_tensor(v).dtype,
                               device=torch.as_tensor(v).device) for k, v in experience.items()}
        dev = next(iter(exps.values())).device
        return BufferState(exps, torch.zeros((), dtype=torch.int64, device=dev), torch.zeros((), dtype=torch.int64, device=dev))

    def add(self, state: BufferState, experience: Dict[str, torch.Tensor]) -> BufferState:
        """buffers.py:40-55 (in place)."""
        return self.add_many(state, {k: torch.as_tensor(v, device=state.current_idx.device).unsqueeze(0)
                                     for k, v in experience.items()})

    def add_many(self, state: BufferState, experiences: Dict[str, torch.Tensor]) -> BufferState:
        """buffers.py:57-77: scatter ``k`` experiences at ``(current_idx + arange(k)) % buffer_size``.
        Updates the storage and the counters IN PLACE (graph-capturable) and returns the same state."""
        k = next(iter(experiences.values())).shape[0]
        dev = state.current_idx.device
        # More experiences than slots: the ring indices repeat, and the reference's ``.at[indices].set`` keeps the LAST
        # write per slot -- i.e. only the last ``buffer_size`` experiences survive, at the same slots.  index_copy_ with
        # duplicate indices is nondeterministic, so write exactly those.
        first = max(k - self.buffer_size, 0)
        idx = (state.current_idx + torch.arange(first, k, device=dev)) % self.buffer_size
        for name, field in state.experiences.items():
            field.index_copy_(0, idx, experiences[name][first:].to(field.dtype))
        state.current_idx.copy_((state.current_idx + k) % self.buffer_size)
        state.current_size.copy_(torch.clamp(state.current_size + k, max=self.buffer_size))
        return state

    def sample(self, state: BufferState, generator=None) -> Dict[str, torch.Tensor]:
        """buffers.py:79-90: ``sample_batch_size`` uniform indices in ``[0, current_size)`` (with replacement)."""
        dev = state.current_idx.device
        # integer draws (the reference uses jax.random.randint): 63 random bits modulo the size -- every slot reachable
        # at any buffer size, modulo bias < 2^-40; no float32 rounding (floor(rand_f32 * size) cannot reach all slots
        # beyond 2^24 entries and is biased below that)
        bits = torch.empty(self.sample_batch_size, dtype=torch.int64, device=dev).random_(generator=generator)
        idx = bits % torch.clamp(state.current_size, min=1)
        return {k: v.index_select(0, idx) for k, v in state.experiences.items()}

    def can_sample(self, state: BufferState) -> bool:
        """buffers.py:92-93 (host-side read of the device counter)."""
        return int(state.current_size.item()) >= self.sample_batch_size
