"""ctypes binding of libb2ntt.so and the reference-shaped host classes on top of it."""
import ctypes as C
import enum
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LAYOUT_COMPACT, LAYOUT_BITSLICED = 0, 1

_lib = None


class B2NTTError(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        msg = lib().b2ntt_strerror(code).decode()
        cuda = lib().b2ntt_last_cuda_error().decode()
        super().__init__(f"{where}: {msg} ({code})" + (f" [{cuda}]" if cuda and code == -3 else ""))


def lib_path():
    # B2NTT_LIB: development override to compare differently tuned builds of the same library
    return os.environ.get("B2NTT_LIB") or os.path.join(HERE, "libb2ntt.so")


def lib():
    """The product library.  Raises if it was not built: there is no other compute path."""
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise ImportError(f"{p} is missing: build it with `make -C binius-ntt_b200/csrc` (or __graft_entry__.build())")
        L = C.CDLL(p)
        vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
        L.b2ntt_plan_create.argtypes = [C.POINTER(vp), i, i, i, i, i, i]
        L.b2ntt_plan_destroy.argtypes = [vp]
        L.b2ntt_exec.argtypes = [vp, vp, vp, vp]
        L.b2ntt_exec_inverse.argtypes = [vp, vp, vp, i, vp]
        L.b2ntt_exec_host.argtypes = [vp, vp, sz, vp]
        L.b2ntt_exec_host_batch.argtypes = [vp, vp, sz, vp, i]
        L.b2ntt_exec_profile.argtypes = [vp, vp, vp, vp, vp, vp, i]
        L.b2ntt_exec_coset.argtypes = [vp, vp, vp, i, vp]
        L.b2ntt_exec_top.argtypes = [vp, vp, i, vp]
        L.b2ntt_plan_launches.argtypes = [vp]
        L.b2ntt_plan_passes.argtypes = [vp]
        L.b2ntt_plan_constants.argtypes = [vp, vp]
        L.b2ntt_bitslice.argtypes = [vp, sz, i, i, vp]
        L.b2ntt_mul.argtypes = [i, i, vp, vp, vp, sz, vp]
        L.b2ntt_mul_repeat.argtypes = [i, vp, sz, i, vp]
        L.b2ntt_check_gpu_capabilities.argtypes = [i]
        L.b2ntt_sumcheck_create.argtypes = [C.POINTER(vp), i, i, i]
        L.b2ntt_sumcheck_destroy.argtypes = [vp]
        L.b2ntt_sumcheck_load_host.argtypes = [vp, vp, i]
        L.b2ntt_sumcheck_load_device.argtypes = [vp, vp, i]
        L.b2ntt_sumcheck_round_messages.argtypes = [vp, vp, vp]
        L.b2ntt_sumcheck_next_round.argtypes = [vp, vp]
        L.b2ntt_sumcheck_round.argtypes = [vp]
        L.b2ntt_sumcheck_launches.argtypes = [vp]
        L.b2ntt_sumcheck_set_recompute.argtypes = [vp, i]
        L.b2ntt_sumcheck_round_static.argtypes = [vp, i, i, vp, vp, vp]
        L.b2ntt_general_create.argtypes = [C.POINTER(vp), i, vp, i]
        L.b2ntt_general_destroy.argtypes = [vp]
        L.b2ntt_general_exec.argtypes = [vp, vp, vp, i, vp]
        L.b2ntt_general_constants.argtypes = [vp, vp]
        L.b2ntt_fourstep_create.argtypes = [C.POINTER(vp), i, i, i, i, i]
        L.b2ntt_fourstep_destroy.argtypes = [vp]
        L.b2ntt_fourstep_block_bytes.argtypes = [vp]
        L.b2ntt_fourstep_block_bytes.restype = sz
        L.b2ntt_fourstep_launches.argtypes = [vp]
        L.b2ntt_fourstep_set_peer.argtypes = [vp, i, vp]
        L.b2ntt_fourstep_scatter.argtypes = [vp, vp, vp]
        L.b2ntt_fourstep_local.argtypes = [vp, vp, vp]
        L.b2ntt_ipc_export.argtypes = [vp, vp, C.POINTER(sz)]
        L.b2ntt_ipc_open.argtypes = [vp, sz, i, C.POINTER(vp)]
        L.b2ntt_ipc_close.argtypes = [vp, sz, i]
        L.b2ntt_mg_create.argtypes = [C.POINTER(vp), i, i, C.POINTER(i), i, i, i, i]
        L.b2ntt_mg_columns.argtypes = [vp, i, C.POINTER(i), C.POINTER(i)]
        L.b2ntt_mg_exec.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
        L.b2ntt_mg_destroy.argtypes = [vp]
        L.b2ntt_bb31_create.argtypes = [C.POINTER(vp), C.c_uint32, i, i, i]
        L.b2ntt_bb31_destroy.argtypes = [vp]
        L.b2ntt_bb31_exec.argtypes = [vp, vp, vp, i, vp]
        L.b2ntt_bb31_exec_host.argtypes = [vp, vp, sz, i, vp]
        L.b2ntt_bb31_launches.argtypes = [vp]
        L.b2ntt_sumcheck_launches.restype = C.c_longlong
        L.b2ntt_strerror.argtypes = [i]
        L.b2ntt_strerror.restype = C.c_char_p
        L.b2ntt_last_cuda_error.restype = C.c_char_p
        _lib = L
    return _lib


def _check(rc, where):
    if rc != 0:
        raise B2NTTError(rc, where)


def _stream_ptr(stream):
    if stream is None:
        import torch

        return C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if isinstance(stream, int):
        return C.c_void_p(stream)
    return C.c_void_p(stream.cuda_stream)


class Plan:
    """b2ntt_plan: domain constants + pass schedule for (field, layout, log_h, log_rate, batch)."""

    def __init__(self, field_bits, log_h, log_rate=0, layout=LAYOUT_COMPACT, batch=1, device=0):
        self._h = C.c_void_p()
        _check(lib().b2ntt_plan_create(C.byref(self._h), field_bits, layout, log_h, log_rate, batch, device), "b2ntt_plan_create")
        self.field_bits, self.log_h, self.log_rate, self.layout, self.batch, self.device = field_bits, log_h, log_rate, layout, batch, device
        self.limbs = field_bits // 32
        self.in_words = (batch << log_h) * self.limbs
        self.out_words = self.in_words << log_rate

    def close(self):
        if self._h:
            lib().b2ntt_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launches(self):
        return lib().b2ntt_plan_launches(self._h)

    @property
    def passes(self):
        return lib().b2ntt_plan_passes(self._h)

    def constants(self):
        out = np.zeros((32, 32), dtype=np.uint32)
        _check(lib().b2ntt_plan_constants(self._h, out.ctypes.data), "b2ntt_plan_constants")
        return out

    def _dev_ok(self, t, words, name):
        import torch

        if not (t.is_cuda and t.is_contiguous() and t.dtype in (torch.int32, torch.uint32)):
            raise ValueError(f"{name}: need a contiguous CUDA int32/uint32 tensor")
        if t.numel() != words:
            raise ValueError(f"{name}: expected {words} words, got {t.numel()}")
        if t.device.index != self.device:
            raise ValueError(f"{name}: tensor lives on cuda:{t.device.index}, the plan on cuda:{self.device}")

    def exec(self, d_in, d_out=None, stream=None):
        """Forward transform of device tensors (32-bit words).  Returns d_out."""
        import torch

        self._dev_ok(d_in, self.in_words, "d_in")
        if d_out is None:
            d_out = d_in if self.log_rate == 0 else torch.empty(self.out_words, dtype=d_in.dtype, device=d_in.device)
        self._dev_ok(d_out, self.out_words, "d_out")
        _check(lib().b2ntt_exec(self._h, d_in.data_ptr(), d_out.data_ptr(), _stream_ptr(stream)), "b2ntt_exec")
        return d_out

    def exec_profile(self, d_in, d_out=None, stream=None):
        """exec with per-pass device times: returns [(kernel name, ms), ...] (synchronises)."""
        self._dev_ok(d_in, self.in_words, "d_in")
        if d_out is None:
            d_out = d_in
        self._dev_ok(d_out, self.out_words, "d_out")
        ms = (C.c_float * 16)()
        kind = (C.c_int * 16)()
        n = lib().b2ntt_exec_profile(self._h, d_in.data_ptr(), d_out.data_ptr(), _stream_ptr(stream), ms, kind, 16)
        _check(min(n, 0), "b2ntt_exec_profile")
        return [(("b2::pass_kernel", "b2::upper_kernel", "b2::bottom_kernel", "b2::upper_kernel", "b2::bsl_kernel")[kind[i]], float(ms[i])) for i in range(n)]

    def exec_inverse(self, d_in, d_out=None, coset=0, stream=None):
        self._dev_ok(d_in, self.in_words, "d_in")
        if d_out is None:
            d_out = d_in
        self._dev_ok(d_out, self.in_words, "d_out")
        _check(lib().b2ntt_exec_inverse(self._h, d_in.data_ptr(), d_out.data_ptr(), coset, _stream_ptr(stream)), "b2ntt_exec_inverse")
        return d_out

    def exec_coset(self, d_in, d_out=None, coset=0, stream=None):
        """Forward transform of ONE coset row (row `coset` of exec's output); in place by default."""
        self._dev_ok(d_in, self.in_words, "d_in")
        if d_out is None:
            d_out = d_in
        self._dev_ok(d_out, self.in_words, "d_out")
        _check(lib().b2ntt_exec_coset(self._h, d_in.data_ptr(), d_out.data_ptr(), coset, _stream_ptr(stream)), "b2ntt_exec_coset")
        return d_out

    def exec_top(self, d_buf, log_segments, stream=None):
        """Top `log_segments` stages of this (full) transform on one device's segments, in place."""
        self._dev_ok(d_buf, self.in_words >> log_segments, "d_buf")
        _check(lib().b2ntt_exec_top(self._h, d_buf.data_ptr(), log_segments, _stream_ptr(stream)), "b2ntt_exec_top")
        return d_buf

    def exec_host(self, h_in, h_out=None):
        """Host numpy (or pinned torch CPU tensor) in -> host out: H2D + transform + D2H + sync."""
        a = _host_words(h_in)
        if h_out is None:
            h_out = np.empty(self.out_words, dtype=np.uint32)
        o = _host_words(h_out)
        if o.size != self.out_words:
            raise ValueError("h_out: wrong size")
        n_in = a.size // self.limbs // self.batch if a.size % (self.limbs * self.batch) == 0 else 0
        _check(lib().b2ntt_exec_host(self._h, a.ctypes.data, n_in, o.ctypes.data), "b2ntt_exec_host")
        return h_out


def _exec_host_batch(self, h_ins, h_outs):
    """Pipelined host path: len(h_ins) independent transforms, H2D / kernels / D2H overlapped on three streams."""
    ins = [_host_words(x) for x in h_ins]
    outs = [_host_words(x) for x in h_outs]
    if len(ins) != len(outs) or any(o.size != self.out_words for o in outs) or any(a.size != self.in_words for a in ins):
        raise ValueError("exec_host_batch: wrong sizes")
    n = len(ins)
    pin = (C.c_void_p * n)(*[a.ctypes.data for a in ins])
    pout = (C.c_void_p * n)(*[o.ctypes.data for o in outs])
    _check(lib().b2ntt_exec_host_batch(self._h, pin, self.in_words // self.limbs // self.batch, pout, n), "b2ntt_exec_host_batch")
    return h_outs


Plan.exec_host_batch = _exec_host_batch


def _host_words(x):
    if isinstance(x, np.ndarray):
        if x.dtype != np.uint32 or not x.flags.c_contiguous:
            raise ValueError("host buffers must be C-contiguous uint32")
        return x.reshape(-1)
    # torch CPU tensor (possibly pinned): view it through numpy without copying
    return x.view(-1).numpy().view(np.uint32)


def bitslice_(d_buf, width_bits, inverse=False, stream=None):
    """In-place compact <-> bit-sliced conversion of a device tensor (transpose_kernel<W> /
    untranspose_kernel<W>, utils/bitslicing.cuh:89-104)."""
    n_blocks, rem = divmod(d_buf.numel(), width_bits)
    if rem:
        raise ValueError("buffer is not a whole number of 32-element blocks")
    _check(lib().b2ntt_bitslice(d_buf.data_ptr(), n_blocks, width_bits, 1 if inverse else 0, _stream_ptr(stream)), "b2ntt_bitslice")
    return d_buf


def mul(a, b, field_bits, layout=LAYOUT_COMPACT, out=None, stream=None):
    """Element-wise tower product of device tensors of 32-bit words."""
    import torch

    if out is None:
        out = torch.empty_like(a)
    words_per = 4 if field_bits == 128 else 1
    n = a.numel() // words_per
    _check(lib().b2ntt_mul(field_bits, layout, a.data_ptr(), b.data_ptr(), out.data_ptr(), n, _stream_ptr(stream)), "b2ntt_mul")
    return out


def mul_repeat(d_io, field_bits, iters, stream=None):
    n_threads = d_io.numel() // field_bits
    _check(lib().b2ntt_mul_repeat(field_bits, d_io.data_ptr(), n_threads, iters, _stream_ptr(stream)), "b2ntt_mul_repeat")
    return d_io


def check_gpu_capabilities(device=0):
    rc = lib().b2ntt_check_gpu_capabilities(device)
    if rc < 0:
        raise B2NTTError(rc, "b2ntt_check_gpu_capabilities")
    return bool(rc)


# ---- the reference's own class names (ntt/nttconf.cuh, ntt/additive_ntt.cuh) -------------------------
class DataOrder(enum.Enum):
    IN_ORDER = 0
    BIT_REVERSED = 1


class NTTData:
    """NTTData<E> (ntt/nttconf.cuh:11-21): host buffer + size + order tag."""

    def __init__(self, order=DataOrder.IN_ORDER, size=0, limbs=1):
        self.order, self.size, self.limbs = order, size, limbs
        self.data = np.zeros(size * limbs, dtype=np.uint32)


class AdditiveNTTConf:
    """AdditiveNTTConf<T,P>(log_h, log_rate) (ntt/nttconf.cuh:49-60).  field_bits 32 = (uint32_t,
    FanPaarTowerField<5>), 128 = GF(2^128) with four limbs."""

    def __init__(self, log_h, log_rate, field_bits=32):
        if not (log_h >= 1 and 0 <= log_rate <= 4 and log_h + log_rate <= 32):
            raise ValueError("AdditiveNTTConf: need 1 <= log_h, 0 <= log_rate <= 4, log_h + log_rate <= 32")
        self.log_h, self.log_rate, self.field_bits = log_h, log_rate, field_bits


class AdditiveNTT:
    """AdditiveNTT<T,P>(conf) with bool apply(in, out) (ntt/additive_ntt.cuh:178-265)."""

    def __init__(self, conf, device=0):
        self.conf = conf
        self.plan = Plan(conf.field_bits, conf.log_h, conf.log_rate, device=device)

    def apply(self, inp, out):
        if inp.size != (1 << self.conf.log_h) or inp.order != DataOrder.IN_ORDER:
            return False  # additive_ntt.cuh:206-208
        n_out = (1 << (self.conf.log_h + self.conf.log_rate)) * self.plan.limbs
        if out.data.size != n_out:
            out.data = np.empty(n_out, dtype=np.uint32)
        out.size = 1 << (self.conf.log_h + self.conf.log_rate)
        out.order = DataOrder.IN_ORDER
        self.plan.exec_host(inp.data, out.data)
        return True


class GeneralNTT:
    """b2ntt_general: GF(2^128) additive NTT on a caller-supplied basis (generalises
    AdditiveNTT<T,P>, ntt/additive_ntt.cuh:176-309, whose basis is fixed to 1 << i).  basis: [log_h, 4] uint32 or None."""

    def __init__(self, log_h, basis=None, device=0):
        self._h = C.c_void_p()
        ptr = None
        if basis is not None:
            self._basis = np.ascontiguousarray(basis, dtype=np.uint32).reshape(-1)
            if self._basis.size != 4 * log_h:
                raise ValueError("basis: expected log_h x 4 words")
            ptr = self._basis.ctypes.data
        _check(lib().b2ntt_general_create(C.byref(self._h), log_h, ptr, device), "b2ntt_general_create")
        self.log_h, self.device, self.words = log_h, device, 4 << log_h

    def close(self):
        if self._h:
            lib().b2ntt_general_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def constants(self):
        out = np.zeros((self.log_h, self.log_h, 4), dtype=np.uint32)
        _check(lib().b2ntt_general_constants(self._h, out.ctypes.data), "b2ntt_general_constants")
        return out

    def exec(self, d_in, d_out=None, inverse=False, stream=None):
        Plan._dev_ok(self, d_in, self.words, "d_in")
        if d_out is None:
            d_out = d_in
        Plan._dev_ok(self, d_out, self.words, "d_out")
        _check(lib().b2ntt_general_exec(self._h, d_in.data_ptr(), d_out.data_ptr(), int(inverse), _stream_ptr(stream)), "b2ntt_general_exec")
        return d_out


MG_COLUMN_SHARD, MG_FOUR_STEP = 0, 1


class MultiGPU:
    """b2ntt_mg: ONE process driving several GPUs of the box.  mode MG_COLUMN_SHARD: `batch` independent columns
    in contiguous slabs per device, no traffic between devices; mode MG_FOUR_STEP: one transform of 2^log_h elements split over
    the devices, the exchange fused into the first phase's peer stores.  exec takes one device tensor per GPU and blocks."""

    def __init__(self, mode, devices, field_bits, log_h, log_rate=0, batch=1):
        self._h = C.c_void_p()
        devs = (C.c_int * len(devices))(*devices)
        _check(lib().b2ntt_mg_create(C.byref(self._h), mode, len(devices), devs, field_bits, log_h, log_rate, batch), "b2ntt_mg_create")
        self.mode, self.devices, self.limbs = mode, list(devices), field_bits // 32

    def columns(self, i):
        first, count = C.c_int(), C.c_int()
        _check(lib().b2ntt_mg_columns(self._h, i, C.byref(first), C.byref(count)), "b2ntt_mg_columns")
        return first.value, count.value

    def exec(self, d_ins, d_outs):
        n = len(self.devices)
        if len(d_ins) != n or len(d_outs) != n:
            raise ValueError("MultiGPU.exec: one input and one output tensor per device")
        for t, d in zip(list(d_ins) + list(d_outs), self.devices * 2):
            if t is not None and (not t.is_cuda or t.device.index != d or not t.is_contiguous()):
                raise ValueError("MultiGPU.exec: tensor i must be contiguous on devices[i]")
        pin = (C.c_void_p * n)(*[t.data_ptr() if t is not None else None for t in d_ins])
        pout = (C.c_void_p * n)(*[t.data_ptr() if t is not None else None for t in d_outs])
        _check(lib().b2ntt_mg_exec(self._h, pin, pout), "b2ntt_mg_exec")
        return d_outs

    def close(self):
        if self._h:
            lib().b2ntt_mg_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


BB31_P = 15 * (1 << 27) + 1
_BB31_R = (1 << 32) % BB31_P
_BB31_RINV = pow(_BB31_R, BB31_P - 2, BB31_P)


def bb31_encode(x):
    """canonical residues (any uint32 is reduced first, as BB31(uint32_t) does) -> risc0::Fp raw words (Montgomery form)"""
    return ((np.asarray(x, dtype=np.uint64) % BB31_P) * _BB31_R % BB31_P).astype(np.uint32)


def bb31_decode(raw):
    """risc0::Fp raw words -> canonical residues (Fp::asUInt32)"""
    return (np.asarray(raw, dtype=np.uint64) * _BB31_RINV % BB31_P).astype(np.uint32)


class BabyBearNTT:
    """NTT<BB31> with NTTConfRad2<BB31>(generator, log_group_order, log_inp_size) (ntt/gpuntt.cuh:126-209,
    ntt/nttconf.cuh:25-47) over b2ntt_bb31_*.  `apply` takes and returns CANONICAL residues as numpy uint32 (the C ABI and
    `exec` move risc0::Fp raw words, the memory image of a BB31 array)."""

    def __init__(self, log_inp_size, generator=137, log_group_order=27, device=0):
        self._h = C.c_void_p()
        gen_raw = int(bb31_encode(np.array([generator], dtype=np.uint32))[0])
        _check(lib().b2ntt_bb31_create(C.byref(self._h), gen_raw, log_group_order, log_inp_size, device), "b2ntt_bb31_create")
        self.log_n, self.device, self.in_words = log_inp_size, device, 1 << log_inp_size

    def close(self):
        if self._h:
            lib().b2ntt_bb31_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launches(self):
        return lib().b2ntt_bb31_launches(self._h)

    def apply(self, x, bit_reversed=False):
        raw = np.ascontiguousarray(bb31_encode(x))
        out = np.empty_like(raw)
        _check(lib().b2ntt_bb31_exec_host(self._h, raw.ctypes.data, raw.size, int(bit_reversed), out.ctypes.data), "b2ntt_bb31_exec_host")
        return bb31_decode(out)

    def exec(self, d_in, d_out=None, bit_reversed=False, stream=None):
        """device tensors of raw words; out of place unless the input is already bit-reversed"""
        import torch

        Plan._dev_ok(self, d_in, self.in_words, "d_in")
        if d_out is None:
            d_out = d_in if bit_reversed else torch.empty_like(d_in)
        Plan._dev_ok(self, d_out, self.in_words, "d_out")
        _check(lib().b2ntt_bb31_exec(self._h, d_in.data_ptr(), d_out.data_ptr(), int(bit_reversed), _stream_ptr(stream)), "b2ntt_bb31_exec")
        return d_out


class Sumcheck:
    """Sumcheck<NUM_VARS, COMPOSITION_SIZE, DATA_IS_TRANSPOSED> (sumcheck/sumcheck.cuh:10-300): prover for the sum over the
    hypercube of a product of `composition_size` multilinears over GF(2^128).  `evals`: composition_size columns of
    2^num_vars evaluations as 32-bit words, compact or (data_is_transposed) bit-sliced in batches of 32 -- a host numpy
    array (copied, as the reference's constructor does) or a CUDA tensor (adopted and overwritten in place)."""

    def __init__(self, evals, num_vars, composition_size, data_is_transposed=False, device=None):
        import torch

        if device is None:
            device = evals.device.index if isinstance(evals, torch.Tensor) and evals.is_cuda else torch.cuda.current_device()
        self.num_vars, self.composition_size = num_vars, composition_size
        self._h = C.c_void_p()
        _check(lib().b2ntt_sumcheck_create(C.byref(self._h), num_vars, composition_size, device), "b2ntt_sumcheck_create")
        words = 4 * composition_size << num_vars
        if isinstance(evals, torch.Tensor) and evals.is_cuda:
            if evals.numel() != words or evals.element_size() != 4 or not evals.is_contiguous():
                raise ValueError("Sumcheck: evals must be a contiguous tensor of composition_size * 2^num_vars * 4 32-bit words")
            self._keep = evals
            _check(lib().b2ntt_sumcheck_load_device(self._h, evals.data_ptr(), int(data_is_transposed)), "b2ntt_sumcheck_load_device")
        else:
            import numpy as np

            h = np.ascontiguousarray(evals, dtype=np.uint32).reshape(-1)
            if h.size != words:
                raise ValueError("Sumcheck: evals must hold composition_size * 2^num_vars * 4 32-bit words")
            _check(lib().b2ntt_sumcheck_load_host(self._h, h.ctypes.data, int(data_is_transposed)), "b2ntt_sumcheck_load_host")

    def set_recompute(self, on=True):
        """compute p(1) from the data in every device round and fail (B2NTT_ECHECK) if it differs from the derived value"""
        _check(lib().b2ntt_sumcheck_set_recompute(self._h, int(on)), "b2ntt_sumcheck_set_recompute")

    def this_round_messages(self):
        """-> (sum[4], points[composition_size + 1][4]) as uint32 arrays"""
        import numpy as np

        s = np.zeros(4, dtype=np.uint32)
        p = np.zeros(4 * (self.composition_size + 1), dtype=np.uint32)
        _check(lib().b2ntt_sumcheck_round_messages(self._h, s.ctypes.data, p.ctypes.data), "b2ntt_sumcheck_round_messages")
        return s, p.reshape(-1, 4)

    def move_to_next_round(self, challenge):
        import numpy as np

        c = np.ascontiguousarray(challenge, dtype=np.uint32).reshape(-1)
        if c.size != 4:
            raise ValueError("Sumcheck: a challenge is 4 32-bit words")
        _check(lib().b2ntt_sumcheck_next_round(self._h, c.ctypes.data), "b2ntt_sumcheck_next_round")

    @property
    def round(self):
        return lib().b2ntt_sumcheck_round(self._h)

    @property
    def launches(self):
        return lib().b2ntt_sumcheck_launches(self._h)

    def close(self):
        if getattr(self, "_h", None):
            lib().b2ntt_sumcheck_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
