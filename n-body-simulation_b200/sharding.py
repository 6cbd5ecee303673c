"""Host-side arithmetic of the multi-GPU layout: who owns which particles and where they sit in
the replicated position array.  Pure Python (no CUDA): the same numbers the C library derives in
``nbody_cuda_configure``; unit-tested on CPU and, with ``gloo``, across processes.

Sharding is the reference's: contiguous index slices, counts[r] = n/P + (r < n%P)
(src/apps/headless_main.cpp:249-259), each rank accumulating full rows for its slice
(src/MpiNaiveSimulation.cpp:51-102).  What differs is the exchange: every rank keeps ALL positions
(16 B per particle) and one all-gather -- or direct NVLink stores -- republishes the slice each
step, instead of rotating 40-byte blocks round a ring (src/MpiNaiveSimulation.cpp:106-120).
"""
from __future__ import annotations

from dataclasses import dataclass

SLOT_ALIGN = 256   # csrc/nbody_kernels.cuh: kSlotAlign


@dataclass(frozen=True)
class SlotLayout:
    """Replicated position array: ``world * slot_stride`` float4 slots; rank r's particle l lives in
    slot ``r * slot_stride + l``; slots past ``counts[r]`` are zero-mass padding.  Equal strides make
    the all-gather a plain equal-count, in-place collective even when n % world != 0."""

    n_total: int
    world: int

    @property
    def counts(self):
        base, rem = divmod(self.n_total, self.world)
        return [base + (1 if r < rem else 0) for r in range(self.world)]

    @property
    def displs(self):
        out, acc = [], 0
        for c in self.counts:
            out.append(acc)
            acc += c
        return out

    @property
    def slot_stride(self) -> int:
        m = max(max(self.counts), 1)
        return (m + SLOT_ALIGN - 1) // SLOT_ALIGN * SLOT_ALIGN

    @property
    def n_slots(self) -> int:
        return self.slot_stride * self.world

    def slot_of(self, i: int) -> int:
        """Slot of global particle i."""
        d = self.displs
        for r in range(self.world - 1, -1, -1):
            if i >= d[r] and self.counts[r] > 0:
                return r * self.slot_stride + (i - d[r])
        raise IndexError(i)

    def owner_of(self, i: int) -> int:
        return self.slot_of(i) // self.slot_stride


def allgather_slices(full, rank: int, layout: SlotLayout, dist, group=None, floats_per_slot: int = 4):
    """In-place all-gather of a replicated slot array: rank r contributes its own stride-sized
    segment of ``full`` (a 1-D tensor of ``n_slots * floats_per_slot`` elements)."""
    per = layout.slot_stride * floats_per_slot
    mine = full[rank * per:(rank + 1) * per]
    dist.all_gather_into_tensor(full, mine, group=group)
    return full
