"""The C-ABI boundary without a GPU: the library loads, exports every symbol include/caribou_cuda.h declares
(and nothing is declared that is not exported), struct layouts of the ctypes mirror match the header's, and
compute entry points fail loudly - not silently on a CPU path - when no device is present."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "caribou_cuda.h")


def _declared_functions():
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(caribou_cuda_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_are_exported(cb):
    names = _declared_functions()
    assert len(names) >= 30
    lib = C.CDLL(cb.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert sorted(cb.EXPORTED_SYMBOLS) == names


def test_no_torch_or_python_in_the_abi(cb):
    """The boundary is plain C: the shared object must not depend on torch / python libraries."""
    out = subprocess.run(["ldd", cb.LIB_PATH], capture_output=True, text=True).stdout
    assert "torch" not in out and "python" not in out, out


def test_struct_layouts_match_the_header(cb, tmp_path):
    """sizeof/offsetof of the ctypes mirrors against a C translation unit that includes the real header."""
    src = tmp_path / "layout.c"
    src.write_text('''
#include <stdio.h>
#include <stddef.h>
#include "caribou_cuda.h"
int main(void) {
  printf("%zu %zu %zu %zu %zu %zu\\n", sizeof(caribou_cuda_config), offsetof(caribou_cuda_config, sos),
         offsetof(caribou_cuda_config, cuda_stream), offsetof(caribou_cuda_config, kernel_form),
         sizeof(caribou_cuda_unpack_out), sizeof(caribou_cuda_result));
  printf("%zu %zu %zu\\n", sizeof(caribou_cuda_filter_info_t), offsetof(caribou_cuda_filter_info_t, decay_samples),
         offsetof(caribou_cuda_filter_info_t, note));
  printf("%zu %zu\\n", sizeof(caribou_cuda_sample_complex_int16), sizeof(caribou_cuda_sample_meta));
  return 0; }''')
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    a, b, c = subprocess.check_output([str(exe)], text=True).strip().split("\n")
    assert [int(v) for v in a.split()] == [C.sizeof(cb.Config), cb.Config.sos.offset, cb.Config.cuda_stream.offset,
                                           cb.Config.kernel_form.offset, C.sizeof(cb.UnpackOut), C.sizeof(cb.Result)]
    assert [int(v) for v in b.split()] == [C.sizeof(cb.FilterInfo), cb.FilterInfo.decay_samples.offset,
                                           cb.FilterInfo.note.offset]
    assert [int(v) for v in c.split()] == [4, 1]          # packed, as caribou_smi.h:53-64


def test_host_side_design_needs_no_gpu(cb, po):
    """caribou_cuda_butter_lowpass is pure host code: same bits as the oracle's restatement of iir1."""
    for order, fc in [(6, 10e3), (6, 25e3), (6, 50e3), (4, 250e3), (5, 100e3), (2, 1e6)]:
        np.testing.assert_array_equal(cb.butter_lowpass(order, 4e6, fc), po.butter_lowpass_sos(order, 4e6, fc))


def test_compute_fails_loudly_without_a_gpu(cb):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(cb.CaribouCudaError) as e:
        cb.Context()
    assert "no CUDA device" in str(e.value) or "CUDA" in str(e.value)
    lib = cb.lib()
    assert lib.caribou_cuda_smi_open(0, 0, 0) is None or True   # must not crash; a device is needed later


def test_tone_tables_of_the_package_are_the_generators(cb, po):
    """bench.py and the multi-GPU scripts take the synthetic generator's tone tables from the package (the B200 arm
    never imports the oracle); they must be the tables the CPU generator uses."""
    a, b = cb.tone_tables()
    c, d = po.tone_tables()
    np.testing.assert_array_equal(a, c)
    np.testing.assert_array_equal(b, d)


def test_stream_result_layout_and_mem_io(cb, tmp_path):
    src = tmp_path / "layout2.c"
    src.write_text('''
#include <stdio.h>
#include <stddef.h>
#include "caribou_cuda.h"
int main(void) {
  printf("%zu %zu %d %d\\n", sizeof(caribou_cuda_stream_result), offsetof(caribou_cuda_stream_result, first_bad_chunk),
         CARIBOU_CUDA_MEM_IO(CARIBOU_CUDA_MEM_HOST_PINNED, CARIBOU_CUDA_MEM_HOST), CARIBOU_CUDA_SHARD_ID_BYTES);
  return 0; }''')
    exe = tmp_path / "layout2"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    v = [int(x) for x in subprocess.check_output([str(exe)], text=True).split()]
    assert v == [C.sizeof(cb.StreamResult), cb.StreamResult.first_bad_chunk.offset,
                 cb.MEM_IO(cb.MEM_HOST_PINNED, cb.MEM_HOST), cb.SHARD_ID_BYTES]


def test_sharding_needs_a_context_and_fails_loudly(cb):
    """No GPU: the sharding entry points reject a null context instead of crashing; the NCCL binding is lazy, so
    the library loads where libnccl is absent."""
    L = cb.lib()
    assert L.caribou_cuda_shard_init(None, 0, 1, None) == cb.ERR
    assert L.caribou_cuda_shard_submit(None, 0, None, 0, 0, None, 0) == cb.ERR
    assert L.caribou_cuda_shard_finish(None) == cb.ERR
    assert L.caribou_cuda_stream_results(None, None) == cb.ERR
