"""Row-band data parallelism across the GPUs of one box.

Every pixel is independent; the only cross-pixel coupling in the reference is the RNG seed = global pixel id
(raytracing.cl:33) and id -> (x, y) (raygen.cl:13-14), both preserved by rendering rows at their global ids.
Bands of `band_rows` rows are dealt out cyclically (band b -> rank b % world) so that every rank gets the
same mix of cheap (sky/floor) and expensive (glass/mirror) rows; each rank renders its bands into a COMPACT
buffer and ONE collective per frame -- a gather to rank 0 -- followed by one interleaving copy rebuilds the
row-major frame.  No other communication exists on this path.

Works on any torch.distributed backend: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
import torch
import torch.distributed as dist


def band_rows_of(rank, world, H, band_rows):
    """Global row indices owned by `rank`, in compact (local) order.  Python twin of rtb_partition_rows."""
    rows = []
    nb = (H + band_rows - 1) // band_rows
    for b in range(rank, nb, world):
        rows.extend(range(b * band_rows, min(H, (b + 1) * band_rows)))
    return rows


def max_rows(world, H, band_rows):
    return max(len(band_rows_of(r, world, H, band_rows)) for r in range(world))


class FrameGather:
    """Pre-allocated gather of compact band buffers into rank 0's row-major frame."""

    def __init__(self, W, H, band_rows, world, rank, device, dtype=torch.int32):
        self.W, self.H, self.band, self.world, self.rank = W, H, band_rows, world, rank
        self.max_rows = max_rows(world, H, band_rows)
        self.local = torch.zeros(self.max_rows * W, dtype=dtype, device=device)  # this rank's compact buffer
        self.uniform = H % (band_rows * world) == 0
        if rank == 0:
            self.parts = [torch.empty_like(self.local) for _ in range(world)]
            self.frame = torch.empty(H * W, dtype=dtype, device=device)
            if not self.uniform:  # ragged tail: precomputed gather index (row granularity)
                src = torch.empty(H, dtype=torch.long)
                for r in range(world):
                    for lr, y in enumerate(band_rows_of(r, world, H, band_rows)):
                        src[y] = r * self.max_rows + lr
                self.src_rows = src.to(device)
        else:
            self.parts, self.frame = None, None

    def gather(self):
        """The one collective of the frame.  Returns the frame on rank 0, None elsewhere."""
        if self.world == 1:
            self.frame.copy_(self.local[: self.H * self.W])
            return self.frame
        dist.gather(self.local, self.parts, dst=0)
        if self.rank != 0:
            return None
        W, nb = self.W, (self.H + self.band - 1) // self.band
        if self.uniform:
            g = torch.stack(self.parts).view(self.world, nb // self.world, self.band * W)
            self.frame.view(nb // self.world, self.world, self.band * W).copy_(g.permute(1, 0, 2))
        else:
            allrows = torch.stack(self.parts).view(self.world * self.max_rows, W)
            torch.index_select(allrows, 0, self.src_rows, out=self.frame.view(self.H, W))
        return self.frame
