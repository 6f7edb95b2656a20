"""ctypes bindings for the CPU oracle (oracle/liboracle.so) and, when present, the
reference's own code (oracle/_ref/libcaribou_ref.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Never imported by cariboulite_b200.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_SO = os.path.join(HERE, "_ref", "libcaribou_ref.so")

NATIVE_BATCH_LEN = 524288
CH_S1G, CH_HIF = 0, 1
FILT_NONE, FILT_20K, FILT_50K, FILT_100K = 0, 1, 2, 3   # CaribouliteStream.hpp:77-84
# Stream filter -> (order, fs, fc) exactly as CaribouliteStream.cpp:85-91 sets them up
STREAM_FILTERS = {FILT_20K: (6, 4e6, 20e3 / 2), FILT_50K: (6, 4e6, 50e3 / 2), FILT_100K: (6, 4e6, 100e3 / 2)}


def build(ref=True):
    """Compile the oracle (and oracle/_ref when /root/reference exists)."""
    subprocess.check_call(["make", "-s", "-C", HERE, "oracle"])
    if ref:
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


_p = C.c_void_p
_u8p = C.POINTER(C.c_uint8)


def _arr(a, dtype):
    a = np.ascontiguousarray(a, dtype=dtype)
    return a


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            build(ref=False)
        L = C.CDLL(ORACLE_SO)
        L.orc_find_buffer_offset.restype = C.c_int
        L.orc_find_buffer_offset.argtypes = [_p, C.c_size_t]
        L.orc_rx_data_analyze.restype = C.c_int
        L.orc_rx_data_analyze.argtypes = [C.c_int, _p, C.c_size_t, _p, _p]
        L.orc_smi_read.restype = C.c_long
        L.orc_smi_read.argtypes = [_p, C.c_size_t, C.POINTER(C.c_size_t), C.c_size_t, C.c_int, _p, _p, C.c_size_t]
        L.orc_flag_scan.restype = None
        L.orc_flag_scan.argtypes = [_p, C.c_size_t, _p, C.c_size_t, C.POINTER(C.c_uint64), _p, C.c_size_t,
                                    C.POINTER(C.c_uint64)]
        L.orc_demux_gather.restype = C.c_size_t
        L.orc_demux_gather.argtypes = [_p, C.c_size_t, C.c_size_t, C.c_int, _p]
        L.orc_gen_words.restype = None
        L.orc_gen_words.argtypes = [C.c_uint64, C.c_size_t, C.c_int, C.c_uint64, _p, _p, _p]
        L.orc_butter_lowpass_sos.restype = C.c_int
        L.orc_butter_lowpass_sos.argtypes = [C.c_int, C.c_double, C.c_double, _p]
        L.orc_cascade_create.restype = _p
        L.orc_cascade_create.argtypes = [C.c_int, _p]
        L.orc_cascade_destroy.argtypes = [_p]
        L.orc_cascade_reset.argtypes = [_p]
        L.orc_cascade_get_state.argtypes = [_p, _p]
        L.orc_cascade_set_state.argtypes = [_p, _p]
        L.orc_cascade_run_f32.argtypes = [_p, _p, _p, C.c_size_t]
        L.orc_cascade_run_f64.argtypes = [_p, _p, _p, C.c_size_t]
        L.orc_cascade_run_cs16_inplace.argtypes = [_p, _p, _p, C.c_size_t]
        L.orc_rx_path_f32.restype = C.c_long
        L.orc_rx_path_f32.argtypes = [_p, C.c_size_t, C.c_size_t, C.c_int, _p, _p, C.c_int, _p, C.c_size_t]
        for f in ("orc_cs16_to_cf32", "orc_cs16_to_cf64", "orc_cs16_to_cs8"):
            getattr(L, f).argtypes = [_p, _p, C.c_size_t]
            getattr(L, f).restype = None
        for f in ("orc_tx_pack", "orc_tx_from_cf32", "orc_tx_from_cf64", "orc_tx_from_cs8"):
            getattr(L, f).argtypes = [_p, C.c_size_t, _p]
            getattr(L, f).restype = None
        _lib = L
    return _lib


# ------------------------------------------------------------------ synthetic input
def tone_tables():
    """One 800-sample period of round(3000*cos(2*pi*5e3*n/4e6) + 800*cos(2*pi*1e6*n/4e6)) (I rail)
    and the same with sin (Q rail).  5 kHz has period 800 and 1 MHz period 4 at fs = 4 MHz, so the
    table makes host and device generators bit-identical without device transcendentals."""
    n = np.arange(800, dtype=np.float64)
    ti = np.round(3000.0 * np.cos(2 * np.pi * n / 800.0) + 800.0 * np.cos(2 * np.pi * n / 4.0))
    tq = np.round(3000.0 * np.sin(2 * np.pi * n / 800.0) + 800.0 * np.sin(2 * np.pi * n / 4.0))
    return ti.astype(np.int16), tq.astype(np.int16)


def gen_words(n, n0=0, channel=CH_S1G, seed=None):
    """Synthetic SMI words; seed defaults to 0xC0FFEE + channel."""
    if seed is None:
        seed = 0xC0FFEE + channel
    ti, tq = tone_tables()
    out = np.empty(n, dtype=np.uint32)
    lib().orc_gen_words(n0, n, int(channel == CH_HIF), seed, _ptr(ti), _ptr(tq), _ptr(out))
    return out


def make_word(i, q, s=0, channel=CH_S1G):
    """SMI word with a valid header carrying 13-bit two's complement I,Q (A0)."""
    i = np.asarray(i, dtype=np.int64) & 0x1FFF
    q = np.asarray(q, dtype=np.int64) & 0x1FFF
    s = np.asarray(s, dtype=np.int64) & 1
    hi, lo = (i, q) if channel == CH_S1G else (q, i)
    return (0x80004000 | (hi << 17) | (lo << 1) | s).astype(np.uint32)


# ------------------------------------------------------------------ unpack stage
def find_buffer_offset(buf):
    b = _arr(buf, np.uint8)
    return lib().orc_find_buffer_offset(_ptr(b), b.size)


def smi_read(raw, length_samples, channel=CH_S1G, native_batch_len=NATIVE_BATCH_LEN, cursor=0,
             cs16=None, meta=None, want_cs16=True, want_meta=True, sentinel=None):
    """orc_smi_read over a byte stream. Returns (ret, cs16[n,2] int16, meta[n] uint8, new_cursor)."""
    raw = np.ascontiguousarray(raw).view(np.uint8).reshape(-1)
    if cs16 is None and want_cs16:
        cs16 = np.zeros((length_samples + 1, 2), dtype=np.int16)
        if sentinel is not None:
            cs16[:] = sentinel
    if meta is None and want_meta:
        meta = np.zeros(length_samples + 1, dtype=np.uint8)
        if sentinel is not None:
            meta[:] = sentinel & 0xFF
    cur = C.c_size_t(cursor)
    ret = lib().orc_smi_read(_ptr(raw), raw.size, C.byref(cur), native_batch_len, channel,
                             _ptr(cs16), _ptr(meta), length_samples)
    return ret, cs16, meta, cur.value


def flag_scan(words, cap=None):
    w = _arr(words, np.uint32)
    cap = w.size if cap is None else cap
    inv = np.zeros(cap, dtype=np.uint32)
    syn = np.zeros(cap, dtype=np.uint32)
    ni, ns = C.c_uint64(0), C.c_uint64(0)
    lib().orc_flag_scan(_ptr(w), w.size, _ptr(inv), cap, C.byref(ni), _ptr(syn), cap, C.byref(ns))
    return ni.value, inv[:min(ni.value, cap)].copy(), ns.value, syn[:min(ns.value, cap)].copy()


def demux_gather(words, B, sub):
    w = _arr(words, np.uint32)
    out = np.empty(w.size, dtype=np.uint32)
    m = lib().orc_demux_gather(_ptr(w), w.size, B, sub, _ptr(out))
    return out[:m].copy()


def cs16_to_cf32(iq):
    iq = _arr(iq, np.int16)
    out = np.empty(iq.shape, dtype=np.float32)
    lib().orc_cs16_to_cf32(_ptr(iq), _ptr(out), iq.size // 2)
    return out


def cs16_to_cf64(iq):
    iq = _arr(iq, np.int16)
    out = np.empty(iq.shape, dtype=np.float64)
    lib().orc_cs16_to_cf64(_ptr(iq), _ptr(out), iq.size // 2)
    return out


def cs16_to_cs8(iq):
    iq = _arr(iq, np.int16)
    out = np.empty(iq.shape, dtype=np.int8)
    lib().orc_cs16_to_cs8(_ptr(iq), _ptr(out), iq.size // 2)
    return out


# ------------------------------------------------------------------ TX direction
def tx_pack(iq):
    """caribou_smi_generate_data (caribou_smi.c:688-717) on real samples: CS16 pairs -> 32-bit TX words."""
    iq = _arr(iq, np.int16)
    out = np.empty(iq.size // 2, dtype=np.uint32)
    lib().orc_tx_pack(_ptr(iq), iq.size // 2, _ptr(out))
    return out


def tx_to_cs16(x, fmt):
    """Stream::WriteSamples conversions (CaribouliteStream.cpp:199-244): CF32 / CF64 / CS8 -> native int16 pairs."""
    dt, fn = {"CF32": (np.float32, "orc_tx_from_cf32"), "CF64": (np.float64, "orc_tx_from_cf64"),
              "CS8": (np.int8, "orc_tx_from_cs8")}[fmt]
    x = _arr(x, dt)
    out = np.empty(x.shape, dtype=np.int16)
    getattr(lib(), fn)(_ptr(x), x.size // 2, _ptr(out))
    return out


# ------------------------------------------------------------------ IIR
def butter_lowpass_sos(order, fs, fc):
    sos = np.zeros(((order + 1) // 2, 6), dtype=np.float64)
    k = lib().orc_butter_lowpass_sos(order, fs, fc, _ptr(sos))
    if k < 0:
        raise ValueError("bad Butterworth parameters")
    return sos


class Cascade:
    """CascadeStages<K, DirectFormII> restatement (one rail)."""

    def __init__(self, sos):
        self.sos = _arr(sos, np.float64).reshape(-1, 6)
        self.k = self.sos.shape[0]
        self.h = lib().orc_cascade_create(self.k, _ptr(self.sos))

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_cascade_destroy(self.h)
            self.h = None

    def reset(self):
        lib().orc_cascade_reset(self.h)

    @property
    def state(self):
        v = np.zeros(2 * self.k)
        lib().orc_cascade_get_state(self.h, _ptr(v))
        return v

    @state.setter
    def state(self, v):
        v = _arr(v, np.float64)
        lib().orc_cascade_set_state(self.h, _ptr(v))

    def run_f32(self, x):
        x = _arr(x, np.float32)
        y = np.empty_like(x)
        lib().orc_cascade_run_f32(self.h, _ptr(x), _ptr(y), x.size)
        return y

    def run_f64(self, x):
        x = _arr(x, np.float64)
        y = np.empty_like(x)
        lib().orc_cascade_run_f64(self.h, _ptr(x), _ptr(y), x.size)
        return y


def filter_cs16_q16(sos, iq, ci=None, cq=None):
    """Reference ordering (A6 'Q16'): in-place int16 filter, returns the filtered CS16 copy."""
    iq = np.array(iq, dtype=np.int16, copy=True).reshape(-1, 2)
    ci = ci or Cascade(sos)
    cq = cq or Cascade(sos)
    lib().orc_cascade_run_cs16_inplace(ci.h, cq.h, _ptr(iq), iq.shape[0])
    return iq


def filter_cf32_f32(sos, x, ci=None, cq=None):
    """North-star ordering (A6 'F32'): x is CF32 [n,2]; filter<float> per rail."""
    x = _arr(x, np.float32).reshape(-1, 2)
    ci = ci or Cascade(sos)
    cq = cq or Cascade(sos)
    y = np.empty_like(x)
    y[:, 0] = ci.run_f32(np.ascontiguousarray(x[:, 0]))
    y[:, 1] = cq.run_f32(np.ascontiguousarray(x[:, 1]))
    return y


class StreamOracle:
    """Restatement of SoapySDR::Stream's RX half (LC/soapy_api/CaribouliteStream.cpp:51-97,
    127-141, 260-382): persistent MTU-sized intermediate buffer, six never-reset filters,
    MTU clamp for non-CS16 formats, Q16 ordering."""

    def __init__(self, raw, channel=CH_S1G, native_batch_len=NATIVE_BATCH_LEN):
        self.raw = np.ascontiguousarray(raw).view(np.uint8).reshape(-1)
        self.channel = channel
        self.nbl = native_batch_len
        self.cursor = 0
        self.mtu = native_batch_len // 4
        self.interm = np.zeros((self.mtu + 1, 2), dtype=np.int16)
        self.fmt = "CS16"
        self.filt = FILT_NONE
        self.filters = {}
        for t, (order, fs, fc) in STREAM_FILTERS.items():
            sos = butter_lowpass_sos(order, fs, fc)
            self.filters[t] = (Cascade(sos), Cascade(sos))

    def set_format(self, fmt):
        if fmt not in ("CS16", "CS8", "CF32", "CF64"):
            return -1
        self.fmt = fmt
        return 0

    def set_filter(self, t):
        self.filt = t if t in self.filters else FILT_NONE

    def _read_cs16(self, buf, n):
        ret, _, _, self.cursor = smi_read(self.raw, n, self.channel, self.nbl, self.cursor, cs16=buf,
                                          want_meta=False)
        if ret < 0:
            ret = 0                                        # CaribouliteStream.cpp:266-276
        if self.filt != FILT_NONE and ret > 0:
            ci, cq = self.filters[self.filt]
            lib().orc_cascade_run_cs16_inplace(ci.h, cq.h, _ptr(buf), ret)
        return ret

    def read(self, n):
        """ReadSamplesGen: returns (ret, array) in the current format."""
        if self.fmt == "CS16":
            buf = np.zeros((n + 1, 2), dtype=np.int16)
            ret = self._read_cs16(buf, n)
            return ret, buf[:max(ret, 0)]
        n = min(n, self.mtu)
        ret = self._read_cs16(self.interm, n)
        src = self.interm[:ret]
        if self.fmt == "CF32":
            return ret, cs16_to_cf32(src)
        if self.fmt == "CF64":
            return ret, cs16_to_cf64(src)
        return ret, cs16_to_cs8(src)


# ------------------------------------------------------------------ the reference itself
_ref = None


def ref_available():
    return os.path.exists(REF_SO)


def ref():
    global _ref
    if _ref is None:
        R = C.CDLL(REF_SO)
        R.ref_dev_open.restype = _p
        R.ref_dev_open.argtypes = [_p, C.c_size_t, C.c_size_t, C.c_int]
        R.ref_dev_close.argtypes = [_p]
        R.ref_dev_rewind.argtypes = [_p]
        for f in ("ref_dev_smi_read", "ref_dev_radio_read"):
            getattr(R, f).restype = C.c_int
            getattr(R, f).argtypes = [_p, _p, _p, C.c_size_t]
        R.ref_dev_read_only.restype = C.c_long
        R.ref_dev_read_only.argtypes = [_p, C.c_size_t]
        R.ref_dev_smi_write.restype = C.c_int
        R.ref_dev_smi_write.argtypes = [_p, _p, C.c_size_t]
        R.ref_dev_read_back.restype = C.c_long
        R.ref_dev_read_back.argtypes = [_p, _p, C.c_size_t]
        R.ref_ring_create.restype = _p
        R.ref_ring_create.argtypes = [C.c_size_t, C.c_int, C.c_int]
        R.ref_ring_destroy.argtypes = [_p]
        R.ref_ring_put.restype = C.c_size_t
        R.ref_ring_put.argtypes = [_p, _p, C.c_size_t]
        R.ref_ring_get.restype = C.c_size_t
        R.ref_ring_get.argtypes = [_p, _p, C.c_size_t, C.c_int]
        for f in ("ref_ring_size", "ref_ring_capacity"):
            getattr(R, f).restype = C.c_size_t
            getattr(R, f).argtypes = [_p]
        R.ref_ring_reset.argtypes = [_p]
        R.ref_stream_open.argtypes = [_p]
        R.ref_stream_set_format.argtypes = [_p, C.c_char_p]
        R.ref_stream_set_format.restype = C.c_int
        R.ref_stream_set_filter.argtypes = [_p, C.c_int]
        R.ref_stream_mtu.restype = C.c_size_t
        R.ref_stream_mtu.argtypes = [_p]
        R.ref_stream_read.restype = C.c_int
        R.ref_stream_read.argtypes = [_p, _p, C.c_size_t]
        R.ref_butter_create.restype = _p
        R.ref_butter_create.argtypes = [C.c_int, C.c_double, C.c_double]
        R.ref_sos_create.restype = _p
        R.ref_sos_create.argtypes = [C.c_int, _p]
        R.ref_filt_destroy.argtypes = [_p]
        R.ref_filt_reset.argtypes = [_p]
        R.ref_filt_sos.restype = C.c_int
        R.ref_filt_sos.argtypes = [_p, _p]
        R.ref_filt_run_f32.argtypes = [_p, _p, _p, C.c_size_t]
        R.ref_filt_run_f64.argtypes = [_p, _p, _p, C.c_size_t]
        R.ref_filt_run_cs16_inplace.argtypes = [_p, _p, _p, C.c_size_t]
        R.ref_rx_path_f32.restype = C.c_long
        R.ref_rx_path_f32.argtypes = [_p, _p, _p, C.c_int, _p, C.c_size_t]
        _ref = R
    return _ref


class RefDev:
    """The reference's caribou_smi_read / Stream on a memfd holding `raw`."""

    def __init__(self, raw, channel=CH_S1G, native_batch_len=NATIVE_BATCH_LEN):
        raw = np.ascontiguousarray(raw).view(np.uint8).reshape(-1)
        self.h = ref().ref_dev_open(_ptr(raw), raw.size, native_batch_len, channel)
        if not self.h:
            raise RuntimeError("ref_dev_open failed")
        self.fmt = "CS16"

    def close(self):
        if self.h:
            ref().ref_dev_close(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def smi_read(self, n, sentinel=None, want_cs16=True, want_meta=True):
        cs16 = np.zeros((n + 1, 2), dtype=np.int16) if want_cs16 else None
        meta = np.zeros(n + 1, dtype=np.uint8) if want_meta else None
        if sentinel is not None:
            if cs16 is not None:
                cs16[:] = sentinel
            if meta is not None:
                meta[:] = sentinel & 0xFF
        ret = ref().ref_dev_smi_read(self.h, _ptr(cs16), _ptr(meta), n)
        return ret, cs16, meta

    def rewind(self):
        ref().ref_dev_rewind(self.h)

    def read_only(self, n):
        """The read(2) loop of caribou_smi_read without the decode (what the memfd copy costs)."""
        return ref().ref_dev_read_only(self.h, n)

    def smi_read_into(self, cs16, meta, n):
        """caribou_smi_read into caller-owned arrays (no allocation in the timed region)."""
        return ref().ref_dev_smi_read(self.h, _ptr(cs16), _ptr(meta), n)

    def stream_read_into(self, buf, n):
        return ref().ref_stream_read(self.h, _ptr(buf), n)

    def smi_write(self, cs16):
        """The reference's caribou_smi_write on the memfd: (return value, the 32-bit words it handed to the fd)."""
        cs16 = _arr(cs16, np.int16)
        n = cs16.size // 2
        ret = ref().ref_dev_smi_write(self.h, _ptr(cs16), n)
        out = np.zeros(n * 4 + 16, dtype=np.uint8)
        got = ref().ref_dev_read_back(self.h, _ptr(out), out.size)
        return ret, out[:got].view(np.uint32).copy() if got % 4 == 0 else out[:got].copy()

    def stream_open(self):
        ref().ref_stream_open(self.h)

    def stream_set_format(self, fmt):
        r = ref().ref_stream_set_format(self.h, fmt.encode())
        if r == 0:
            self.fmt = fmt
        return r

    def stream_set_filter(self, t):
        ref().ref_stream_set_filter(self.h, t)

    def stream_read(self, n):
        dt = {"CS16": np.int16, "CS8": np.int8, "CF32": np.float32, "CF64": np.float64}[self.fmt]
        buf = np.zeros((n + 1, 2), dtype=dt)
        ret = ref().ref_stream_read(self.h, _ptr(buf), n)
        return ret, buf[:max(ret, 0)]


class RefFilter:
    """iir1 filter object from the reference build (one rail)."""

    def __init__(self, order=None, fs=None, fc=None, sos=None):
        if sos is not None:
            s = _arr(sos, np.float64).reshape(-1, 6)
            self.h = ref().ref_sos_create(s.shape[0], _ptr(s))
        else:
            self.h = ref().ref_butter_create(order, fs, fc)
        if not self.h:
            raise ValueError("reference filter construction failed")

    def __del__(self):
        if getattr(self, "h", None):
            ref().ref_filt_destroy(self.h)
            self.h = None

    def sos(self):
        out = np.zeros((8, 6))
        k = ref().ref_filt_sos(self.h, _ptr(out))
        return out[:k].copy()

    def reset(self):
        ref().ref_filt_reset(self.h)

    def run_f32(self, x):
        x = _arr(x, np.float32)
        y = np.empty_like(x)
        ref().ref_filt_run_f32(self.h, _ptr(x), _ptr(y), x.size)
        return y

    def run_f64(self, x):
        x = _arr(x, np.float64)
        y = np.empty_like(x)
        ref().ref_filt_run_f64(self.h, _ptr(x), _ptr(y), x.size)
        return y


# ------------------------------------------------------------------ whole RX path (CPU baseline legs)
def rx_path_f32(raw, sos, D, channel=CH_S1G, native_batch_len=NATIVE_BATCH_LEN):
    """Oracle port: raw SMI bytes -> CF32 filtered, decimated [m,2] (one capture, zero initial state)."""
    raw = np.ascontiguousarray(raw).view(np.uint8).reshape(-1)
    n = raw.size // 4
    ci, cq = Cascade(sos), Cascade(sos)
    out = np.zeros((n // D + 1, 2), dtype=np.float32)
    m = lib().orc_rx_path_f32(_ptr(raw), raw.size, native_batch_len, channel, ci.h, cq.h, D, _ptr(out), n)
    if m < 0:
        raise RuntimeError("orc_rx_path_f32 -> %d" % m)
    return out[:m]


class RefRxPath:
    """The reference build's whole path for one capture: memfd device + two iir1 filter objects."""

    def __init__(self, raw, order, fs, fc, D, channel=CH_S1G, native_batch_len=NATIVE_BATCH_LEN):
        self.dev = RefDev(raw, channel, native_batch_len)
        self.fi, self.fq = RefFilter(order, fs, fc), RefFilter(order, fs, fc)
        self.D = D
        self.n = np.ascontiguousarray(raw).view(np.uint8).size // 4
        self.out = np.zeros((self.n // D + 1, 2), dtype=np.float32)

    def rewind(self):
        ref().ref_dev_rewind(self.dev.h)
        self.fi.reset()
        self.fq.reset()

    def run(self):
        m = ref().ref_rx_path_f32(self.dev.h, self.fi.h, self.fq.h, self.D, _ptr(self.out), self.n)
        if m < 0:
            raise RuntimeError("ref_rx_path_f32 -> %d" % m)
        return self.out[:m]
