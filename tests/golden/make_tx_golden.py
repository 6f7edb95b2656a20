"""Generates tests/golden/tx_golden.npz by running the REFERENCE's own caribou_smi_write (oracle/_ref, built from
the unmodified sources under /root/reference) on a memfd.  As committed, the reference's TX packer ignores its
samples and packs the constant test pattern ii = 0xFFFF, qq = 0 (caribou_smi.c:701-702), so the fixture pins the
word for the sample (i = -1, q = 0), the chunked write loop and the return value.  Run where /root/reference exists:
    python tests/golden/make_tx_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402

po.build(ref=True)
assert po.ref_available()
NBL = 32768
out = dict(nbl=NBL)
for n in (1, 5, 8192, 8192 * 3 + 77):
    dev = po.RefDev(np.zeros(4, dtype=np.uint32), 0, NBL)
    ret, words = dev.smi_write(np.zeros((n, 2), dtype=np.int16))
    out["ret_%d" % n] = ret
    out["words_%d" % n] = words
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tx_golden.npz"), **out)
print("wrote tx_golden.npz", {k: (np.asarray(v).shape, np.unique(v)[:3]) for k, v in out.items()})
