"""Host side of kernel (b) without a GPU: the filter plan (realisation, run length, steady state for the biased
integer input, correction table R, run matrix powers, state maps) emulated run by run on the CPU exactly as the
kernel combines them, against the oracle's sequential FP64 cascade.  This is the arithmetic the scan kernel
distributes over lanes and tiles; what it adds on the GPU (shuffles, look-back) is covered by the -m gpu tests."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIAS_WORDS = 1052672.0          # 2^20 + 2^12, csrc/iir.cuh kBiasWords


@pytest.fixture(scope="module")
def emul(tmp_path_factory):
    out = tmp_path_factory.mktemp("plan") / "plan_emul.so"
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    subprocess.check_call([nvcc, "-O2", "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-o", str(out),
                           os.path.join(ROOT, "tests", "host", "plan_emul.cpp"),
                           os.path.join(ROOT, "cariboulite_b200", "csrc", "tables.cpp")])
    L = C.CDLL(str(out))
    L.emul_plan_info.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    L.emul_run.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_void_p,
                           C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    L.emul_check_powers.restype = C.c_double
    L.emul_check_powers.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int]
    return L


def plan_info(L, sos, D, y32, force_cascade=0):
    info = np.zeros(10)
    assert L.emul_plan_info(len(sos), sos.ctypes.data, D, force_cascade, y32, info.ctypes.data) == 0
    return dict(par=int(info[0]), window=int(info[1]), chained=int(info[2]), decay=int(info[3]), par_err=info[4],
                hn1=info[5], run_len=int(info[6]), kept=int(info[7]), y32=int(info[8]), rerun=int(info[9]))


def test_run_lengths_follow_the_shared_memory_budget(emul, po):
    """csrc/iir.cuh iir_run_len: a multiple of 32 and of D, at most 32 kept outputs per run and a multiple of 4 of
    them; float staging (plain CF32 output) affords the longer runs."""
    for order in (1, 2, 4, 6, 8):
        sos = po.butter_lowpass_sos(order, 4e6, 100e3)
        for D in (1, 2, 4, 8, 16, 32):
            for y32 in (0, 1):
                p = plan_info(emul, sos, D, y32)
                assert p["run_len"] % 32 == 0 and p["run_len"] % D == 0
                assert p["kept"] == p["run_len"] // D and p["kept"] % 4 == 0
                assert p["par"] == 1                              # Butterworth poles are distinct: parallel form
                if p["rerun"]:
                    # full-rate streams (csrc/iir_rerun.cuh): the input is fetched twice instead of staged, so no
                    # stage of kept outputs bounds the run and there is no float staging
                    assert (D == 1 or (D == 2 and len(sos) >= 3)) and p["y32"] == 0
                    assert p["run_len"] == (128 if len(sos) <= 2 else 256)
                else:
                    assert p["kept"] <= 32 and p["y32"] == y32
    assert plan_info(emul, po.butter_lowpass_sos(4, 4e6, 250e3), 1, 1, force_cascade=1)["rerun"] == 0   # cascade form: staged kernel
    sos4 = po.butter_lowpass_sos(4, 4e6, 250e3)
    assert plan_info(emul, sos4, 8, 1)["run_len"] == 160          # the C4 configuration (profiles/r1_sweeps.txt, sweep 18)
    assert plan_info(emul, sos4, 16, 1)["run_len"] == 320
    assert plan_info(emul, po.butter_lowpass_sos(6, 4e6, 25e3), 8, 1)["run_len"] == 160
    assert plan_info(emul, po.butter_lowpass_sos(8, 4e6, 400e3), 8, 1)["run_len"] == 256


@pytest.mark.parametrize("order,fc", [(1, 100e3), (2, 100e3), (3, 50e3), (4, 250e3), (6, 10e3), (6, 50e3), (8, 400e3)])
@pytest.mark.parametrize("D", [1, 8, 32])
@pytest.mark.parametrize("y32", [0, 1])
def test_plan_emulation_matches_the_sequential_cascade(emul, po, order, fc, D, y32):
    sos = po.butter_lowpass_sos(order, 4e6, fc)
    K = len(sos)
    p = plan_info(emul, sos, D, y32)
    n = p["run_len"] * 40
    rng = np.random.default_rng(order * 100 + D)
    x = (np.round(3000 * np.cos(2 * np.pi * 5e3 * np.arange(n) / 4e6)) + rng.integers(-64, 64, n)).astype(np.float64)
    ref_c = po.Cascade(sos)
    st_in = rng.standard_normal(2 * K) * 100                         # a non-trivial incoming state (DF-II coordinates)
    ref_c.state = st_in
    ref = ref_c.run_f64(x)[::D]
    for force_cascade, bias in ((0, BIAS_WORDS), (0, 0.0), (1, 0.0)):
        out = np.zeros(n // D)
        st_out = np.zeros(2 * K)
        r = emul.emul_run(K, sos.ctypes.data, D, force_cascade, y32, bias, 1.0, x.ctypes.data, n, st_in.ctypes.data,
                          out.ctypes.data, st_out.ctypes.data)
        assert r == (0 if force_cascade else 1)
        scale = np.sqrt(np.mean(ref ** 2))
        # FP64 reordering plus, with the bias, the cancellation of 2^20 * H(1): still far inside the 1e-5 budget
        assert np.sqrt(np.mean((out - ref) ** 2)) / scale < (2e-9 if bias else 1e-11), (force_cascade, bias)
        np.testing.assert_allclose(st_out, ref_c.state, rtol=1e-7, atol=1e-7 * np.abs(ref_c.state).max())


def test_run_matrix_powers_and_sparsity(emul, po):
    for order, fc in [(2, 100e3), (4, 250e3), (6, 10e3), (8, 400e3)]:
        sos = po.butter_lowpass_sos(order, 4e6, fc)
        for D in (1, 8, 32):
            for y32 in (0, 1):
                for force_cascade in (0, 1):
                    w = emul.emul_check_powers(len(sos), sos.ctypes.data, D, force_cascade, y32)
                    assert 0 <= w < 1e-12, (order, D, y32, force_cascade, w)


@pytest.mark.parametrize("order,fc", [(2, 100e3), (4, 250e3), (6, 10e3), (6, 25e3), (6, 50e3), (8, 10e3), (8, 400e3)])
@pytest.mark.parametrize("D,words", [(1, 1), (1, 0), (2, 1)])
def test_ap_form_of_the_rerun_kernel_matches_the_sequential_cascade(emul, po, order, fc, D, words):
    """csrc/iir_rerun.cuh, AP phase 3: (1 + q)^2K on the integer samples (earlier samples zero at every run start),
    the sum dropped into a double's mantissa, all-pole cascade from the mapped entry state - against the oracle's
    sequential cascade, full-scale input, non-trivial incoming state."""
    sos = po.butter_lowpass_sos(order, 4e6, fc)
    K = len(sos)
    emul.emul_run_ap.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_size_t, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_void_p]
    p = plan_info(emul, sos, D, 0)
    if not p["rerun"]:
        pytest.skip("staged kernel for this (order, D)")
    n = p["run_len"] * 24
    rng = np.random.default_rng(order * 10 + D)
    full = 4096 if words else 32768
    x = np.clip(np.round(0.7 * full * np.cos(2 * np.pi * 5e3 * np.arange(n) / 4e6)) + rng.integers(-full // 4, full // 4, n),
                -full, full - 1).astype(np.int32)
    x[:64] = -full                                                    # both extremes of the field
    x[64:128] = full - 1
    ref_c = po.Cascade(sos)
    st_in = rng.standard_normal(2 * K) * 100
    ref_c.state = st_in
    ref = ref_c.run_f64(x.astype(np.float64))[::D]
    out = np.zeros(n // D)
    st_out = np.zeros(2 * K)
    info = np.zeros(3)
    r = emul.emul_run_ap(K, sos.ctypes.data, D, words, 1.0, x.ctypes.data, n, st_in.ctypes.data, out.ctypes.data,
                         st_out.ctypes.data, info.ctypes.data)
    assert r == 1 and info[0] == 1 and info[1] == 2 * K and info[2] < 1e-11, (r, info)
    assert np.sqrt(np.mean((out - ref) ** 2)) / np.sqrt(np.mean(ref ** 2)) < 2e-9


def test_ap_form_is_refused_for_other_numerators(emul, po):
    """Zeros elsewhere than z = -1 (here: a scaled b1) keep the parallel-form phase 3."""
    sos = po.butter_lowpass_sos(4, 4e6, 250e3).copy()
    sos[1, 1] *= 0.999
    emul.emul_run_ap.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_void_p, C.c_size_t, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_void_p]
    x = np.zeros(128 * 4, np.int32)
    st = np.zeros(4)
    out = np.zeros(x.size)
    info = np.zeros(3)
    assert emul.emul_run_ap(2, sos.ctypes.data, 1, 1, 1.0, x.ctypes.data, x.size, st.ctypes.data, out.ctypes.data,
                            st.ctypes.data, info.ctypes.data) == 0 and info[0] == 0
