"""cariboulite_b200 - host-side Python mirror of libcaribou_cuda's C ABI (include/caribou_cuda.h).

The product is the shared library; this module only binds it with ctypes so that tests and
bench.py can drive it.  There is no CPU fallback: importing works without a GPU (so the symbol
table can be checked), every compute call needs a B200 and fails loudly otherwise.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

HERE = os.path.dirname(os.path.abspath(__file__))
# CARIBOU_CUDA_LIB: an alternative build of the same library (kernel tuning experiments); never a fallback
LIB_PATH = os.environ.get("CARIBOU_CUDA_LIB") or os.path.join(HERE, "libcaribou_cuda.so")
HEADER_PATH = os.path.join(HERE, "..", "include", "caribou_cuda.h")

OK, ERR, ERR_DEBUG, ERR_SYNC = 0, -1, -2, -3
CH_S1G, CH_HIF = 0, 1
MEM_DEVICE, MEM_HOST_PINNED, MEM_HOST = 0, 1, 2
FMT_CF32, FMT_CS16, FMT_CS8, FMT_CF64 = 0, 1, 2, 3
PARITY_Q16, PARITY_F32 = 0, 1
FORM_AUTO, FORM_CASCADE, FORM_PARALLEL = 0, 1, 2
IN_WORDS, IN_CS16, IN_CF32 = 0, 1, 2
NATIVE_BATCH_LEN = 524288
MAX_SOS = 4
FMT_BYTES = {FMT_CF32: 8, FMT_CS16: 4, FMT_CS8: 2, FMT_CF64: 16}
FMT_DTYPE = {FMT_CF32: np.float32, FMT_CS16: np.int16, FMT_CS8: np.int8, FMT_CF64: np.float64}
IN_BYTES = {IN_WORDS: 4, IN_CS16: 4, IN_CF32: 8}


class CaribouCudaError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [("device", C.c_int), ("channel", C.c_int), ("interleave_block_words", C.c_uint32),
                ("chunk_bytes", C.c_size_t), ("n_streams", C.c_int), ("n_sos", C.c_int),
                ("sos", (C.c_double * 6) * MAX_SOS), ("decimation", C.c_int), ("parity", C.c_int),
                ("out_format", C.c_int), ("cuda_stream", C.c_void_p), ("kernel_form", C.c_int)]


class FilterInfo(C.Structure):
    _fields_ = [("n_sos", C.c_int), ("decimation", C.c_int), ("form", C.c_int), ("run_len", C.c_int),
                ("tile_samples", C.c_int), ("lookback_window", C.c_int), ("chained", C.c_int),
                ("decay_samples", C.c_uint64), ("parallel_form_error", C.c_double), ("note", C.c_char * 96)]


class UnpackOut(C.Structure):
    _fields_ = [("cs16", C.c_void_p), ("meta", C.c_void_p), ("cf32", C.c_void_p),
                ("cs16_b", C.c_void_p), ("meta_b", C.c_void_p), ("cf32_b", C.c_void_p),
                ("invalid_idx", C.c_void_p), ("invalid_cap", C.c_size_t),
                ("sync_idx", C.c_void_p), ("sync_cap", C.c_size_t)]


class StreamResult(C.Structure):
    _fields_ = [("n_samples", C.c_int64), ("n_filtered", C.c_int64), ("first_bad_chunk", C.c_int32),
                ("n_misaligned_chunks", C.c_int32)]


def MEM_IO(mem_in, mem_out):
    """CARIBOU_CUDA_MEM_IO: different memory kinds for the input bytes and the output arrays of unpack_submit."""
    return 0x100 | (mem_in & 0xF) | ((mem_out & 0xF) << 4)


SHARD_ID_BYTES = 128
SHARD_ALLGATHER, SHARD_CHAIN, SHARD_PEER = 0, 1, 2


class Result(C.Structure):
    _fields_ = [("n_samples", C.c_int64), ("n_samples_b", C.c_int64), ("n_filtered", C.c_int64),
                ("n_invalid", C.c_uint64), ("n_sync", C.c_uint64), ("first_bad_chunk", C.c_int32),
                ("n_misaligned_chunks", C.c_int32)]


_lib = None


def lib():
    """Loads libcaribou_cuda.so (built in-tree by build.py).  No fallback of any kind."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CaribouCudaError("libcaribou_cuda.so is missing: run `python -m cariboulite_b200.build` "
                               "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, sz, i, i64, u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_int64, C.c_uint64
    sig = {
        "caribou_cuda_create": (vp, [C.POINTER(Config)]),
        "caribou_cuda_destroy": (None, [vp]),
        "caribou_cuda_last_error": (C.c_char_p, []),
        "caribou_cuda_get_stream": (vp, [vp]),
        "caribou_cuda_wait": (i64, [vp, C.POINTER(Result)]),
        "caribou_cuda_launch_count": (u64, [vp]),
        "caribou_cuda_set_profiling": (i, [vp, i]),
        "caribou_cuda_read_profile": (i, [vp, vp, i]),
        "caribou_cuda_host_alloc": (vp, [sz]),
        "caribou_cuda_host_free": (None, [vp]),
        "caribou_cuda_unpack_submit": (i, [vp, vp, sz, sz, i, C.POINTER(UnpackOut)]),
        "caribou_cuda_filter_submit": (i, [vp, i, vp, sz, sz, i, vp, sz]),
        "caribou_cuda_rx_submit": (i, [vp, vp, sz, sz, i, C.POINTER(UnpackOut), vp, sz]),
        "caribou_cuda_get_state": (i, [vp, i, vp]),
        "caribou_cuda_set_state": (i, [vp, i, vp]),
        "caribou_cuda_reset": (i, [vp]),
        "caribou_cuda_filter_info": (i, [vp, C.POINTER(FilterInfo)]),
        "caribou_cuda_shard_matrix": (i, [vp, u64, vp]),
        "caribou_cuda_fixup_samples": (sz, [vp, sz]),
        "caribou_cuda_filter_fixup": (i, [vp, i, vp, sz, sz, i, vp, sz, vp]),
        "caribou_cuda_butter_lowpass": (i, [i, C.c_double, C.c_double, vp]),
        "caribou_cuda_synth_words": (i, [vp, vp, u64, sz, i, u64, vp, vp]),
        "caribou_cuda_tx_pack_submit": (i, [vp, i, vp, sz, i, vp]),
        "caribou_cuda_smi_write": (i, [vp, i, vp, sz]),
        "caribou_cuda_txstream_write": (i, [vp, i, C.c_char_p, vp, sz]),
        "caribou_cuda_ring_create": (vp, [sz, sz, i, i]),
        "caribou_cuda_ring_destroy": (None, [vp]),
        "caribou_cuda_ring_put": (sz, [vp, vp, sz]),
        "caribou_cuda_ring_get": (sz, [vp, vp, sz, C.c_long]),
        "caribou_cuda_ring_reset": (None, [vp]),
        "caribou_cuda_ring_size": (sz, [vp]),
        "caribou_cuda_ring_capacity": (sz, [vp]),
        "caribou_cuda_ring_dropped": (u64, [vp]),
        "caribou_cuda_async_open": (vp, [vp, i, sz]),
        "caribou_cuda_async_activate": (None, [vp, i]),
        "caribou_cuda_async_read": (sz, [vp, vp, sz, C.c_long]),
        "caribou_cuda_async_idle": (i, [vp]),
        "caribou_cuda_async_ring": (vp, [vp]),
        "caribou_cuda_async_close": (None, [vp]),
        "caribou_cuda_smi_open": (vp, [i, sz, i]),
        "caribou_cuda_smi_close": (None, [vp]),
        "caribou_cuda_smi_read": (i, [vp, i, vp, vp, sz]),
        "caribou_cuda_smi_get_native_batch_samples": (sz, [vp]),
        "caribou_cuda_rxstream_create": (vp, [vp, i]),
        "caribou_cuda_rxstream_destroy": (None, [vp]),
        "caribou_cuda_rxstream_set_format": (i, [vp, C.c_char_p]),
        "caribou_cuda_rxstream_set_filter": (None, [vp, i]),
        "caribou_cuda_rxstream_mtu": (sz, [vp]),
        "caribou_cuda_rxstream_read": (i, [vp, vp, sz, C.c_long]),
        "caribou_cuda_pump": (i64, [vp, i, i, i, sz, i64]),
        "caribou_cuda_stream_results": (i, [vp, vp]),
        "caribou_cuda_radio_read_cf32": (i, [vp, i, vp, vp, sz]),
        "caribou_cuda_shard_unique_id": (i, [vp]),
        "caribou_cuda_shard_init": (i, [vp, i, i, vp]),
        "caribou_cuda_shard_attach_comm": (i, [vp, i, i, vp]),
        "caribou_cuda_shard_set_mode": (i, [vp, i]),
        "caribou_cuda_shard_mode": (i, [vp]),
        "caribou_cuda_shard_submit": (i, [vp, i, vp, sz, sz, vp, sz]),
        "caribou_cuda_shard_begin": (i, [vp, i, vp, sz, sz, vp, sz]),
        "caribou_cuda_shard_gather_buffer": (vp, [vp, C.POINTER(sz)]),
        "caribou_cuda_shard_finish": (i, [vp]),
        "caribou_cuda_shard_exchange_local": (i, [C.POINTER(vp), i]),
        "caribou_cuda_shard_profile": (i, [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
        "caribou_cuda_unpack_noalloc": (i, [i, vp, i, sz, vp, sz, C.POINTER(UnpackOut), vp]),
    }
    for name, (res, args) in sig.items():
        f = getattr(L, name)
        f.restype = res
        f.argtypes = args
    _lib = L
    return L


EXPORTED_SYMBOLS = [
    "caribou_cuda_create", "caribou_cuda_destroy", "caribou_cuda_last_error", "caribou_cuda_get_stream",
    "caribou_cuda_wait", "caribou_cuda_launch_count", "caribou_cuda_set_profiling",
    "caribou_cuda_read_profile", "caribou_cuda_host_alloc", "caribou_cuda_host_free",
    "caribou_cuda_unpack_submit", "caribou_cuda_filter_submit", "caribou_cuda_rx_submit",
    "caribou_cuda_get_state", "caribou_cuda_set_state", "caribou_cuda_reset", "caribou_cuda_filter_info",
    "caribou_cuda_shard_matrix",
    "caribou_cuda_fixup_samples", "caribou_cuda_filter_fixup", "caribou_cuda_butter_lowpass",
    "caribou_cuda_synth_words", "caribou_cuda_smi_open", "caribou_cuda_smi_close", "caribou_cuda_smi_read",
    "caribou_cuda_smi_get_native_batch_samples", "caribou_cuda_rxstream_create",
    "caribou_cuda_rxstream_destroy", "caribou_cuda_rxstream_set_format", "caribou_cuda_rxstream_set_filter",
    "caribou_cuda_rxstream_mtu", "caribou_cuda_rxstream_read", "caribou_cuda_pump",
    "caribou_cuda_tx_pack_submit", "caribou_cuda_smi_write", "caribou_cuda_txstream_write",
    "caribou_cuda_ring_create", "caribou_cuda_ring_destroy", "caribou_cuda_ring_put", "caribou_cuda_ring_get",
    "caribou_cuda_ring_reset", "caribou_cuda_ring_size", "caribou_cuda_ring_capacity", "caribou_cuda_ring_dropped",
    "caribou_cuda_async_open", "caribou_cuda_async_activate", "caribou_cuda_async_read", "caribou_cuda_async_idle",
    "caribou_cuda_async_ring", "caribou_cuda_async_close",
    "caribou_cuda_stream_results", "caribou_cuda_radio_read_cf32",
    "caribou_cuda_shard_unique_id", "caribou_cuda_shard_init", "caribou_cuda_shard_attach_comm",
    "caribou_cuda_shard_set_mode", "caribou_cuda_shard_mode", "caribou_cuda_shard_submit", "caribou_cuda_shard_begin",
    "caribou_cuda_shard_gather_buffer", "caribou_cuda_shard_finish", "caribou_cuda_shard_exchange_local",
    "caribou_cuda_shard_profile", "caribou_cuda_unpack_noalloc",
]


def last_error():
    return lib().caribou_cuda_last_error().decode()


NOALLOC_MAX_BYTES = 64 << 20
NOALLOC_MAX_CHUNKS = 16384


def unpack_noalloc(raw, nbytes, cs16=None, meta=None, cf32=None, invalid_idx=None, sync_idx=None, result=None,
                   device=0, stream=None, channel=0, chunk_bytes=0):
    """caribou_cuda_unpack_noalloc: kernel (a) on device buffers with the library's fixed scratch array - no
    context, no allocation, asynchronous on `stream` (a cudaStream_t as int, None = default stream).  `result` is
    device memory for one caribou_cuda_result (56 bytes) or None."""
    o = UnpackOut()
    o.cs16, o.meta, o.cf32 = _ptr(cs16), _ptr(meta), _ptr(cf32)
    if invalid_idx is not None:
        o.invalid_idx, o.invalid_cap = _ptr(invalid_idx), invalid_idx.numel()
    if sync_idx is not None:
        o.sync_idx, o.sync_cap = _ptr(sync_idx), sync_idx.numel()
    r = lib().caribou_cuda_unpack_noalloc(device, stream, channel, chunk_bytes, _ptr(raw), nbytes, C.byref(o), _ptr(result))
    if r != 0:
        raise CaribouCudaError(last_error())
    return r


def butter_lowpass(order, fs, fc):
    """Host-side Iir::Butterworth::LowPass<order>::setup(fs, fc); rows (b0,b1,b2,a0,a1,a2)."""
    sos = np.zeros(((order + 1) // 2, 6), dtype=np.float64)
    k = lib().caribou_cuda_butter_lowpass(order, fs, fc, sos.ctypes.data)
    if k < 0:
        raise ValueError("bad Butterworth parameters")
    return sos


def _ptr(x):
    """Device pointer of a torch CUDA tensor, host pointer of a numpy array / torch CPU tensor."""
    if x is None:
        return None
    if isinstance(x, int):
        return x
    if isinstance(x, np.ndarray):
        return x.ctypes.data
    return x.data_ptr()      # torch tensor


def _mem_of(x, default=MEM_DEVICE):
    if isinstance(x, np.ndarray):
        return MEM_HOST
    if hasattr(x, "is_cuda"):
        if x.is_cuda:
            return MEM_DEVICE
        return MEM_HOST_PINNED if x.is_pinned() else MEM_HOST
    return default


class Context:
    """One stream of the RX sample path (or a batch of n_streams equal-length captures)."""

    def __init__(self, device=0, channel=CH_S1G, interleave_block_words=0, chunk_bytes=0, n_streams=1,
                 sos=None, decimation=1, parity=PARITY_Q16, out_format=FMT_CF32, cuda_stream=None,
                 kernel_form=FORM_AUTO):
        cfg = Config()
        cfg.device = device
        cfg.channel = channel
        cfg.interleave_block_words = interleave_block_words
        cfg.chunk_bytes = chunk_bytes
        cfg.n_streams = n_streams
        cfg.decimation = decimation
        cfg.parity = parity
        cfg.out_format = out_format
        cfg.cuda_stream = cuda_stream
        cfg.kernel_form = kernel_form
        self.n_sos = 0
        if sos is not None:
            sos = np.ascontiguousarray(sos, dtype=np.float64).reshape(-1, 6)
            if sos.shape[0] > MAX_SOS:
                raise ValueError("at most %d biquads" % MAX_SOS)
            self.n_sos = sos.shape[0]
            for k in range(self.n_sos):
                for j in range(6):
                    cfg.sos[k][j] = sos[k, j]
        cfg.n_sos = self.n_sos
        self.cfg = cfg
        self.decimation = max(1, decimation)
        self.out_format = out_format
        self.n_streams = max(1, n_streams)
        self.h = lib().caribou_cuda_create(C.byref(cfg))
        if not self.h:
            raise CaribouCudaError(last_error())

    def close(self):
        if getattr(self, "h", None):
            lib().caribou_cuda_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc):
        if rc == ERR:
            raise CaribouCudaError(last_error())
        return rc

    @property
    def cuda_stream(self):
        return lib().caribou_cuda_get_stream(self.h)

    @property
    def launch_count(self):
        return lib().caribou_cuda_launch_count(self.h)

    def set_profiling(self, enable=True):
        return self._check(lib().caribou_cuda_set_profiling(self.h, int(enable)))

    def read_profile(self, cap=256):
        """Device milliseconds of every dominant-kernel launch since the last read (waits for the context)."""
        ms = np.zeros(cap, dtype=np.float32)
        k = self._check(lib().caribou_cuda_read_profile(self.h, ms.ctypes.data, cap))
        return ms[:k].astype(np.float64)

    def wait(self):
        r = Result()
        ret = lib().caribou_cuda_wait(self.h, C.byref(r))
        if ret == ERR:
            raise CaribouCudaError(last_error())
        return ret, r

    def stream_results(self):
        """Per-capture results of the last rx_submit (waits for the context)."""
        arr = (StreamResult * self.n_streams)()
        self._check(lib().caribou_cuda_stream_results(self.h, C.cast(arr, C.c_void_p)))
        return list(arr)

    # -- multi-GPU time sharding, hand-off on the device
    def shard_init(self, rank, world, unique_id=None):
        """unique_id: the 128 bytes of shard_unique_id() from rank 0 (ncclCommInitRank inside the library), or None:
        no communicator - move the end vectors yourself between shard_begin and shard_finish."""
        buf = None
        if unique_id is not None:
            assert len(unique_id) == SHARD_ID_BYTES
            buf = C.create_string_buffer(bytes(unique_id), SHARD_ID_BYTES)
        return self._check(lib().caribou_cuda_shard_init(self.h, rank, world, C.cast(buf, C.c_void_p) if buf is not None else None))

    def shard_set_mode(self, mode):
        return self._check(lib().caribou_cuda_shard_set_mode(self.h, mode))

    @property
    def shard_mode(self):
        return self._check(lib().caribou_cuda_shard_mode(self.h))

    def shard_submit(self, in_kind, x, n_samples, out, in_stride=0, out_stride=0):
        return self._check(lib().caribou_cuda_shard_submit(self.h, in_kind, _ptr(x), n_samples, in_stride, _ptr(out), out_stride))

    def shard_begin(self, in_kind, x, n_samples, out, in_stride=0, out_stride=0):
        return self._check(lib().caribou_cuda_shard_begin(self.h, in_kind, _ptr(x), n_samples, in_stride, _ptr(out), out_stride))

    def shard_finish(self):
        return self._check(lib().caribou_cuda_shard_finish(self.h))

    def shard_profile(self):
        """(pass-1 ms, hand-off + fix-up ms) of the last sharded call; needs set_profiling(True)."""
        a, b = C.c_float(0), C.c_float(0)
        self._check(lib().caribou_cuda_shard_profile(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    # -- kernel (a)
    def unpack_submit(self, raw, nbytes=None, stride_bytes=0, cs16=None, meta=None, cf32=None, cs16_b=None,
                      meta_b=None, cf32_b=None, invalid_idx=None, sync_idx=None, mem=None):
        o = UnpackOut()
        o.cs16, o.meta, o.cf32 = _ptr(cs16), _ptr(meta), _ptr(cf32)
        o.cs16_b, o.meta_b, o.cf32_b = _ptr(cs16_b), _ptr(meta_b), _ptr(cf32_b)
        if invalid_idx is not None:
            o.invalid_idx, o.invalid_cap = _ptr(invalid_idx), invalid_idx.numel() if hasattr(invalid_idx, "numel") else invalid_idx.size
        if sync_idx is not None:
            o.sync_idx, o.sync_cap = _ptr(sync_idx), sync_idx.numel() if hasattr(sync_idx, "numel") else sync_idx.size
        if nbytes is None:
            nbytes = raw.numel() * raw.element_size() if hasattr(raw, "numel") else raw.nbytes
        mem = _mem_of(raw) if mem is None else mem
        return self._check(lib().caribou_cuda_unpack_submit(self.h, _ptr(raw), nbytes, stride_bytes, mem, C.byref(o)))

    # -- kernel (b)
    def filter_submit(self, in_kind, x, n_samples, out, in_stride=0, out_stride=0, mem=None):
        mem = _mem_of(x) if mem is None else mem
        return self._check(lib().caribou_cuda_filter_submit(self.h, in_kind, _ptr(x), n_samples, in_stride, mem,
                                                            _ptr(out), out_stride))

    def filter_fixup(self, in_kind, x, n_samples, out, state_in, in_stride=0, out_stride=0):
        st = np.ascontiguousarray(state_in, dtype=np.float64)
        return self._check(lib().caribou_cuda_filter_fixup(self.h, in_kind, _ptr(x), n_samples, in_stride, MEM_DEVICE,
                                                           _ptr(out), out_stride, st.ctypes.data))

    # -- fused
    def rx_submit(self, raw, nbytes, filt_out, stride_bytes=0, filt_stride=0, mem=None, unpack_out=None):
        mem = _mem_of(raw) if mem is None else mem
        uo = C.byref(unpack_out) if unpack_out is not None else None
        return self._check(lib().caribou_cuda_rx_submit(self.h, _ptr(raw), nbytes, stride_bytes, mem, uo,
                                                        _ptr(filt_out), filt_stride))

    # -- state
    def get_state(self, stream=0):
        st = np.zeros(4 * self.n_sos, dtype=np.float64)
        self._check(lib().caribou_cuda_get_state(self.h, stream, st.ctypes.data))
        return st

    def set_state(self, state, stream=0):
        st = np.ascontiguousarray(state, dtype=np.float64)
        assert st.size == 4 * self.n_sos
        self._check(lib().caribou_cuda_set_state(self.h, stream, st.ctypes.data))

    def reset(self):
        self._check(lib().caribou_cuda_reset(self.h))

    def pump(self, in_kind, in_fd, out_fd, block_samples, max_samples=0):
        """Capture pump (file / pipe in, file / pipe out); returns input samples consumed or a negative code."""
        r = lib().caribou_cuda_pump(self.h, in_kind, in_fd, out_fd, block_samples, max_samples)
        if r == ERR:
            raise CaribouCudaError(last_error())
        return r

    def filter_info(self):
        fi = FilterInfo()
        self._check(lib().caribou_cuda_filter_info(self.h, C.byref(fi)))
        return fi

    def shard_matrix(self, n_samples):
        s = 2 * self.n_sos
        m = np.zeros((s, s), dtype=np.float64)
        self._check(lib().caribou_cuda_shard_matrix(self.h, n_samples, m.ctypes.data))
        return m

    def fixup_samples(self, n_samples):
        return lib().caribou_cuda_fixup_samples(self.h, n_samples)

    def synth_words(self, dev_words, n, n0=0, hif_layout=0, seed=0xC0FFEE, tone_i=None, tone_q=None):
        ti = np.ascontiguousarray(tone_i, dtype=np.int16)
        tq = np.ascontiguousarray(tone_q, dtype=np.int16)
        assert ti.size == 800 and tq.size == 800
        return self._check(lib().caribou_cuda_synth_words(self.h, _ptr(dev_words), n0, n, hif_layout, seed,
                                                          ti.ctypes.data, tq.ctypes.data))

    # -- kernel (c): TX direction
    def tx_pack_submit(self, in_format, samples, n, words, mem=None):
        """n samples of `in_format` (FMT_*) -> n SMI TX words; numpy arrays are host memory, torch CUDA tensors
        device memory (include/caribou_cuda.h, caribou_cuda_tx_pack_submit)."""
        if mem is None:
            mem = MEM_HOST if isinstance(samples, np.ndarray) else MEM_DEVICE
        return self._check(lib().caribou_cuda_tx_pack_submit(self.h, in_format, _ptr(samples), n, mem, _ptr(words)))


def shard_unique_id():
    """ncclGetUniqueId through the library (rank 0 calls it and ships the bytes to the other ranks)."""
    buf = C.create_string_buffer(SHARD_ID_BYTES)
    if lib().caribou_cuda_shard_unique_id(C.cast(buf, C.c_void_p)) != OK:
        raise CaribouCudaError(last_error())
    return buf.raw


def shard_exchange_local(contexts):
    """The hand-off between `world` contexts of this process (between shard_begin and shard_finish of all of them)."""
    arr = (C.c_void_p * len(contexts))(*[c.h for c in contexts])
    if lib().caribou_cuda_shard_exchange_local(arr, len(contexts)) != OK:
        raise CaribouCudaError(last_error())


def tone_tables():
    """The 800-entry tone tables of the synthetic SMI word generator: one period of
    round(3000 cos(2 pi 5e3 n / 4e6) + 800 cos(2 pi 1e6 n / 4e6)) for the I rail and the same with sin for the
    Q rail (5 kHz repeats every 800 samples at fs = 4 MHz, 1 MHz every 4).  caribou_cuda_synth_words takes them
    from the host so that the device generator needs no transcendentals; tests/test_abi.py checks them against
    the CPU generator's own tables."""
    n = np.arange(800, dtype=np.float64)
    ti = np.round(3000.0 * np.cos(2 * np.pi * n / 800.0) + 800.0 * np.cos(2 * np.pi * n / 4.0))
    tq = np.round(3000.0 * np.sin(2 * np.pi * n / 800.0) + 800.0 * np.sin(2 * np.pi * n / 4.0))
    return ti.astype(np.int16), tq.astype(np.int16)


def build(force=False):
    return _build.build(force=force)
