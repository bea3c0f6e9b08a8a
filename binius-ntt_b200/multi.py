"""Multi-GPU partitioning of the additive NTT on one NVLink/NVSwitch box.

One process per GPU (`torch.distributed`, backend "nccl"); torch is plumbing only -- the transforms
are libb2ntt.so calls (include/b2ntt.h).

* `ColumnShardedRS` -- config 3: a batch of independent Reed-Solomon column encodings (forward NTT
  with log_rate > 0, every coset of ntt/additive_ntt.cuh:213-215).  Columns never interact, so rank r
  encodes its own contiguous slab of columns; there is NO data-path collective.
* `FourStepNTT`     -- config 4: ONE transform of 2^log_n elements over G = 2^g GPUs.  The reference
  runs its stages from the highest stride down (additive_ntt.cuh:233-247), so the split needs exactly
  one exchange.  Global index i = (h : g bits, m : g bits, l : log_n - 2g bits):
    phase 1  rank m holds the G segments (h, m, .): the top g stages pair elements that differ only
             in h and their twiddles depend only on the h bits above the butterfly bit, so they run
             element-wise across the local segments               -> b2ntt_exec_top
    exchange rank m sends segment h to rank h (all-to-all over NVLink; G-1 of G segments leave)
    phase 2  rank h now owns the contiguous block [h * 2^(log_n-g), ...): an ordinary local
             transform whose twiddle block index is offset by h, i.e. coset h of a
             (log_h = log_n - g, log_rate = g) plan                  -> b2ntt_exec_coset
  Output: natural order, block-sharded (rank h holds block h).
  Two exchange paths: `fused` (default on CUDA with more than one rank) = b2ntt_fourstep_scatter, the phase-1
  kernel stores segment h straight into rank h's block with 16-byte peer stores (blocks mapped across the
  processes by CUDA IPC, two stream-ordered barriers per transform, no collective on the data path);
  otherwise phase 1 in place + NCCL all_to_all_single (the A/B baseline, and gloo on the CPU tests).
"""
import torch
import torch.distributed as dist

from . import api


def shard_columns(n_columns, world, rank):
    """Contiguous slab of columns owned by `rank`: (first, count).  Slabs differ by at most one column."""
    if not (0 <= rank < world) or n_columns < 0:
        raise ValueError("shard_columns: bad rank/world/n_columns")
    base, extra = divmod(n_columns, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def _world_rank(group):
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group), dist.get_rank(group)
    return 1, 0


class ColumnShardedRS:
    """Batched RS encode, columns sharded over ranks (config 3: 256 columns, log_h = 20, log_rate = 2,
    GF(2^32)).  `encode` transforms this rank's columns only; nothing is exchanged."""

    def __init__(self, n_columns, log_h, log_rate, field_bits=32, layout=api.LAYOUT_COMPACT, group=None, device=None):
        self.world, self.rank = _world_rank(group)
        self.first, self.count = shard_columns(n_columns, self.world, self.rank)
        self.n_columns, self.log_h, self.log_rate, self.field_bits = n_columns, log_h, log_rate, field_bits
        self.device = torch.cuda.current_device() if device is None else device
        self.plan = api.Plan(field_bits, log_h, log_rate, layout=layout, batch=self.count, device=self.device) if self.count else None

    @property
    def in_words(self):
        return self.plan.in_words if self.plan else 0

    @property
    def out_words(self):
        return self.plan.out_words if self.plan else 0

    def encode(self, d_in, d_out=None, stream=None):
        """d_in: this rank's `count` columns of 2^log_h elements; returns count * 2^(log_h+log_rate)
        elements, coset-major per column (additive_ntt.cuh:250-259)."""
        if not self.plan:
            return d_in if d_out is None else d_out
        return self.plan.exec(d_in, d_out, stream)

    def encode_from_evaluations(self, d_evals, d_out=None, stream=None):
        """Systematic (interpolate-then-extend) encode, SURVEY 8f.1: d_evals holds this rank's columns as evaluations
        on coset 0; the inverse transform (v ^= u; u ^= w*v, stages low to high) recovers the novel-basis coefficients,
        the forward transform evaluates them on every coset.  Row 0 of each output column equals the input column."""
        if not self.plan:
            return d_evals if d_out is None else d_out
        coeffs = self.plan.exec_inverse(d_evals, torch.empty_like(d_evals), 0, stream)
        return self.plan.exec(coeffs, d_out, stream)

    def close(self):
        if self.plan:
            self.plan.close()
            self.plan = None


class FourStepNTT:
    """One additive NTT of 2^log_n elements split over G = 2^g ranks (config 4: log_n = 30, GF(2^128),
    G = 8: 2 GiB per rank)."""

    def __init__(self, log_n, field_bits=128, layout=api.LAYOUT_COMPACT, group=None, device=None, fused=None):
        self.group = group
        self._fs = None
        self._fused_req = fused
        self.world, self.rank = _world_rank(group)
        g = self.world.bit_length() - 1
        if self.world != 1 << g:
            raise ValueError("FourStepNTT: the number of ranks must be a power of two")
        if layout != api.LAYOUT_COMPACT:
            raise ValueError("FourStepNTT: compact layout only")
        if g > 4:
            raise ValueError("FourStepNTT: at most 16 ranks (coset index bits <= 4, ntt/nttconf.cuh:55-60)")
        self.g, self.log_n, self.field_bits, self.layout = g, log_n, field_bits, layout
        self.limbs = field_bits // 32
        self.log_local = log_n - g                    # elements per rank
        self.seg_words = (1 << (log_n - 2 * g)) * self.limbs if log_n >= 2 * g else 0
        self.local_words = (1 << self.log_local) * self.limbs
        if g and self.seg_words < 32:
            raise ValueError("FourStepNTT: a segment must hold at least one 128-byte line")
        self.device = device
        self._make_plans()

    # -- the two local phases: libb2ntt.so only (the CPU tests override them) ----------------------
    def _make_plans(self):
        dev = torch.cuda.current_device() if self.device is None else self.device
        self.device = dev
        self.full_plan = api.Plan(self.field_bits, self.log_n, 0, layout=self.layout, device=dev) if self.g else None
        self.local_plan = api.Plan(self.field_bits, self.log_local, self.g, layout=self.layout, device=dev)
        want = self._fused_req
        if want is None:
            want = self.g > 0 and dist.is_initialized() and dist.get_backend(self.group) == "nccl"
        if want:
            self._setup_fused(dev)

    def _setup_fused(self, dev):
        """Map every rank's block buffer into this process (CUDA IPC) and create the b2ntt_fourstep object."""
        import ctypes as C

        L = api.lib()
        self._fs = C.c_void_p()
        api._check(L.b2ntt_fourstep_create(C.byref(self._fs), self.field_bits, self.log_n, self.g, self.rank, dev), "b2ntt_fourstep_create")
        self.block = torch.empty(self.local_words, dtype=torch.int32, device=torch.device("cuda", dev))
        handle, off = (C.c_ubyte * 64)(), C.c_size_t()
        api._check(L.b2ntt_ipc_export(self.block.data_ptr(), handle, C.byref(off)), "b2ntt_ipc_export")
        mine = (bytes(handle), off.value)
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=self.group)
        self._opened = []
        for r, (h, o) in enumerate(everyone):
            if r == self.rank:
                ptr = C.c_void_p(self.block.data_ptr())
            else:
                ptr = C.c_void_p()
                api._check(L.b2ntt_ipc_open((C.c_ubyte * 64).from_buffer_copy(h), o, dev, C.byref(ptr)), "b2ntt_ipc_open")
                self._opened.append((ptr, o))
            api._check(L.b2ntt_fourstep_set_peer(self._fs, r, ptr), "b2ntt_fourstep_set_peer")
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.block.device)
        self.exchange_kind = "fused: phase-1 kernel stores into the peers' blocks (16-byte stores over NVLink, CUDA IPC), no collective"

    def _stream_barrier(self):
        """all ranks' work enqueued so far on the current stream is complete before anything enqueued after it starts"""
        dist.all_reduce(self._flag, group=self.group)

    def scatter(self, buf, stream=None):
        """phase 1 + exchange, fused: reads this rank's m-shard, fills every rank's block"""
        api._check(api.lib().b2ntt_fourstep_scatter(self._fs, buf.data_ptr(), api._stream_ptr(stream)), "b2ntt_fourstep_scatter")

    def forward_fused(self, buf):
        """buf: this rank's m-shard (left untouched).  Returns self.block = block `rank` of the transform.  Runs on torch's
        current stream; the two barriers are tiny all-reduces ordered on that stream."""
        if buf.numel() != self.local_words:
            raise ValueError(f"FourStepNTT.forward: expected {self.local_words} words, got {buf.numel()}")
        self._stream_barrier()   # nobody still reads the block of the previous transform
        self.scatter(buf)
        self._stream_barrier()   # every rank's stores have landed
        api._check(api.lib().b2ntt_fourstep_local(self._fs, self.block.data_ptr(), api._stream_ptr(None)), "b2ntt_fourstep_local")
        return self.block

    @property
    def fused(self):
        return self._fs is not None

    @property
    def launches(self):
        """kernel launches of one forward(): top phase + local transform (the exchange is the collective library's)"""
        return (self.full_plan.launches if self.full_plan else 0) + self.local_plan.launches

    exchange_kind = "NCCL all_to_all_single between the two local phases"

    def top(self, buf, stream=None):
        """phase 1 on this rank's G segments (h, m = rank, .), in place."""
        if self.g:
            self.full_plan.exec_top(buf, self.g, stream)
        return buf

    def local(self, buf, stream=None):
        """phase 2 on the block this rank owns, in place."""
        return self.local_plan.exec_coset(buf, buf, self.rank, stream)

    # -- the one exchange ---------------------------------------------------------------------------
    def exchange(self, buf, out=None):
        """segment h of rank m  ->  segment m of rank h.  One all-to-all; `out` must not alias `buf`."""
        if out is None:
            out = torch.empty_like(buf)
        if self.world == 1:
            out.copy_(buf)
            return out
        if dist.get_backend(self.group) != "gloo":
            dist.all_to_all_single(out, buf, group=self.group)  # errors of the collective propagate
        else:
            # gloo has no all-to-all on every build (CPU tests): pairwise sends
            ins = list(buf.view(self.world, -1).unbind(0))
            outs = list(out.view(self.world, -1).unbind(0))
            outs[self.rank].copy_(ins[self.rank])
            reqs = []
            for p in range(self.world):
                if p != self.rank:
                    reqs.append(dist.isend(ins[p].contiguous(), p, group=self.group))
                    reqs.append(dist.irecv(outs[p], p, group=self.group))
            for r in reqs:
                r.wait()
        return out

    def forward(self, buf, scratch=None, stream=None):
        """buf: this rank's m-shard ([h][l] segments, 2^(log_n-g) elements).  Returns the rank's block of
        the transform in natural order (rank h = block h).  The result lives in `scratch` (allocated
        when not given); `buf` is overwritten by phase 1."""
        if buf.numel() != self.local_words:
            raise ValueError(f"FourStepNTT.forward: expected {self.local_words} words, got {buf.numel()}")
        if self._fs is not None and scratch is None and stream is None:
            return self.forward_fused(buf)
        if buf.is_cuda and stream is not None and stream != torch.cuda.current_stream(buf.device):
            # the collective is ordered against torch's CURRENT stream: make `stream` current for all three steps, so that
            # the exchange waits for phase 1, phase 2 waits for the exchange and `scratch` is allocated on the same stream
            with torch.cuda.stream(stream):
                return self.forward(buf, scratch, stream)
        self.top(buf, stream)
        out = self.exchange(buf, scratch)
        return self.local(out, stream)

    # -- host-side helpers for tests and benches ------------------------------------------------------
    @staticmethod
    def m_shard(x_global, world, rank, limbs):
        """This rank's phase-1 input taken from the whole input (any tensor/array with reshape):
        global index (h, m, l) -> local (h, l) for m = rank."""
        if x_global.reshape(-1).shape[0] % (world * world * limbs):
            raise ValueError("m_shard: size is not a multiple of world^2 elements")
        return x_global.reshape(world, world, -1)[:, rank, :].reshape(-1)

    def close(self):
        if self._fs is not None:
            torch.cuda.synchronize()
            if dist.is_initialized():
                dist.barrier(group=self.group)  # no peer still stores into this rank's block
            for ptr, off in self._opened:
                api.lib().b2ntt_ipc_close(ptr, off, self.device)
            api.lib().b2ntt_fourstep_destroy(self._fs)
            self._fs = None
        for p in (getattr(self, "full_plan", None), getattr(self, "local_plan", None)):
            if p:
                p.close()
        self.full_plan = self.local_plan = None
