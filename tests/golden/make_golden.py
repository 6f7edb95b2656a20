"""Generates tests/golden/rx_golden.npz by running the REFERENCE's own code (oracle/_ref, built from the
unmodified sources under /root/reference) on the synthetic input of SURVEY 8(d).  Run it where
/root/reference exists:  python tests/golden/make_golden.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyoracle as po  # noqa: E402

po.build(ref=True)
assert po.ref_available()
N, NBL, DROP = 40000, 32768, 3
w = po.gen_words(N, channel=0)
raw = np.ascontiguousarray(w.view(np.uint8)[DROP:])
dev = po.RefDev(raw, 0, NBL)
ret, cs16, meta = dev.smi_read(N, sentinel=0x1234)
out = dict(n_words=N, nbl=NBL, drop=DROP, words_head=w[:64], ret=ret, cs16=cs16, meta=meta)
for name, (order, fc) in {"sos_10k": (6, 10e3), "sos_25k": (6, 25e3), "sos_50k": (6, 50e3), "sos_c3": (4, 250e3)}.items():
    out[name] = po.RefFilter(order, 4e6, fc).sos()
d2 = po.RefDev(raw, 0, NBL)
d2.stream_open()
d2.stream_set_format("CF32")
d2.stream_set_filter(po.FILT_50K)
reads = [1000, 8192, 5000, 8192]
out["stream_reads"] = np.array(reads)
out["stream_cf32"] = np.concatenate([d2.stream_read(n)[1] for n in reads])
x = po.cs16_to_cf32(cs16[:8192])
fi, fq = po.RefFilter(4, 4e6, 250e3), po.RefFilter(4, 4e6, 250e3)
y = np.stack([fi.run_f32(np.ascontiguousarray(x[:, 0])), fq.run_f32(np.ascontiguousarray(x[:, 1]))], axis=1)
out["f32_c3_d8"] = y[::8]
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "rx_golden.npz"), **out)
print("wrote rx_golden.npz", {k: np.asarray(v).shape for k, v in out.items()})
