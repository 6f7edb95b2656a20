"""Captures of a batch are independent reads (caribou_smi.c:632-682 per capture): one that loses
word alignment must not disturb the others, and submits queued without a wait must not lose a capture that turned
out misaligned (the fused launch is optimistic; the slow path redoes exactly the captures it skipped)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-5
CHUNK = 65536            # bytes: 16384 samples per chunk


def rel_rms(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return np.sqrt(np.mean((a - b) ** 2, axis=0)) / np.sqrt(np.mean(b ** 2, axis=0))


def _dev(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a)).to("cuda:0")


def _oracle_stream(po, sos, D, calls, ci=None, cq=None):
    """What the reference delivers for one capture over several reads: caribou_smi_read into a zeroed buffer, then the
    never-reset filter pair over what the read returned (nothing on -3 beyond the chunks before the bad one)."""
    ci = ci or po.Cascade(sos)
    cq = cq or po.Cascade(sos)
    outs = []
    for w in calls:
        n = w.size
        ret, cs16, _, _ = po.smi_read(w.view(np.uint8), n, 0, CHUNK)
        if ret == -3:
            # the batch API filters the chunks decoded before the un-alignable one (documented in caribou_cuda.h)
            spc = CHUNK // 4
            bad = next(c for c in range(n // spc) if po.find_buffer_offset(w[c * spc:(c + 1) * spc].view(np.uint8)) < 0)
            deliver = bad * spc
        else:
            deliver = n
        deliver -= deliver % D
        outs.append((ret, deliver, po.filter_cf32_f32(sos, po.cs16_to_cf32(cs16[:deliver]), ci, cq)[::D]))
    return outs, np.concatenate([ci.state, cq.state])


@pytest.mark.parametrize("mem", ["device", "pinned"])
def test_one_lost_and_one_misaligned_capture_leave_the_others_alone(cb, po, torch_cuda, mem):
    """Stream 3 of 11 loses sync at chunk 1, stream 7 has a misaligned chunk: the other nine are bit-identical to
    their solo runs on the fused kernel, stream 3 reports -3 like caribou_smi.c:665-668 and delivers chunk 0,
    stream 7 equals the reference's read + filter."""
    torch = torch_cuda
    ns, n, D = 11, 3 * 16384, 8
    spc = CHUNK // 4
    sos = po.butter_lowpass_sos(4, 4e6, 250e3)
    ws = [po.gen_words(n, seed=100 + s) for s in range(ns)]
    ws[3][spc:2 * spc] = 0                           # chunk 1 of stream 3: no valid header anywhere
    ws[7][2 * spc] ^= 0x80000000                     # chunk 2 of stream 7 starts with a bad header (offs = 4)
    raw = np.stack(ws).view(np.int32)
    if mem == "device":
        d_w = _dev(torch, raw)
        out = torch.full((ns, n // D, 2), 7.0, dtype=torch.float32, device="cuda:0")
    else:
        d_w = torch.from_numpy(raw.copy()).pin_memory()
        out = torch.full((ns, n // D, 2), 7.0, dtype=torch.float32).pin_memory()
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32, n_streams=ns, chunk_bytes=CHUNK) as ctx:
        ctx.rx_submit(d_w, n * 4, out, stride_bytes=n * 4, filt_stride=n // D,
                      mem=cb.MEM_DEVICE if mem == "device" else cb.MEM_HOST_PINNED)
        ret, res = ctx.wait()
        sr = ctx.stream_results()
        states = [ctx.get_state(s) for s in range(ns)]
    got = out.cpu().numpy()
    assert ret == cb.ERR_SYNC and res.first_bad_chunk == 1 and res.n_misaligned_chunks == 1
    for s in range(ns):
        (r_ret, deliver, ref), = _oracle_stream(po, sos, D, [ws[s]])[0]
        if s == 3:
            assert sr[s].n_samples == cb.ERR_SYNC and r_ret == -3
            assert sr[s].first_bad_chunk == 1 and sr[s].n_filtered == spc // D == deliver // D
            assert np.all(rel_rms(got[s, :deliver // D], ref) <= TOL)
        else:
            assert sr[s].n_samples == n and sr[s].n_filtered == n // D and sr[s].first_bad_chunk == -1
            assert sr[s].n_misaligned_chunks == (1 if s == 7 else 0)
            assert np.all(rel_rms(got[s], ref) <= TOL), s
    # the nine clean captures: same bits and same end state as when each is submitted alone
    for s in (0, 1, 2, 4, 5, 6, 8, 9, 10):
        solo = torch.zeros((n // D, 2), dtype=torch.float32, device="cuda:0")
        with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32, chunk_bytes=CHUNK) as c1:
            c1.rx_submit(_dev(torch, ws[s].view(np.int32)), n * 4, solo)
            assert c1.wait()[0] == n
            st = c1.get_state()
        assert np.array_equal(got[s], solo.cpu().numpy()), s
        np.testing.assert_array_equal(states[s], st)


@pytest.mark.parametrize("ns", [1, 5])
def test_queued_submits_redo_a_misaligned_capture_in_order(cb, po, torch_cuda, ns):
    """Three rx_submit calls queued before one wait, the FIRST with a misaligned capture: its fused launch is a
    no-op on the device, so the later submits of that capture must not run from the stale state either - all three
    are redone in order at wait(), the other captures stay on the fused kernel."""
    torch = torch_cuda
    n, D, calls = 2 * 16384, 8, 3
    bad = ns // 2
    sos = po.butter_lowpass_sos(6, 4e6, 25e3)
    ws = [[po.gen_words(n, n0=c * n, seed=7 + s) for s in range(ns)] for c in range(calls)]
    ws[0][bad][16384] ^= 0x80000000
    d_w = [_dev(torch, np.stack(ws[c]).view(np.int32)) for c in range(calls)]
    outs = [torch.zeros((ns, n // D, 2), dtype=torch.float32, device="cuda:0") for _ in range(calls)]
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32, n_streams=ns, chunk_bytes=CHUNK) as ctx:
        for c in range(calls):
            ctx.rx_submit(d_w[c], n * 4, outs[c], stride_bytes=n * 4, filt_stride=n // D)
        ret, res = ctx.wait()
        assert ret == n and res.n_misaligned_chunks == 0          # the LAST submit was clean
        states = [ctx.get_state(s) for s in range(ns)]
        # and the context is healthy afterwards: a further submit runs fused again
        extra = torch.zeros((ns, n // D, 2), dtype=torch.float32, device="cuda:0")
        ctx.rx_submit(d_w[1], n * 4, extra, stride_bytes=n * 4, filt_stride=n // D)
        ret2, res2 = ctx.wait()
        assert ret2 == n and res2.n_misaligned_chunks == 0
    for s in range(ns):
        ref, st = _oracle_stream(po, sos, D, [ws[c][s] for c in range(calls)])
        for c in range(calls):
            assert np.all(rel_rms(outs[c][s].cpu().numpy(), ref[c][2]) <= TOL), (s, c)
        np.testing.assert_allclose(states[s], st, rtol=1e-8, atol=1e-9 * np.max(np.abs(st)))


def test_queued_submits_outputs(cb, po, torch_cuda):
    """Same as above with every output checked (separate buffers per submit, pinned host memory)."""
    torch = torch_cuda
    ns, n, D, calls = 6, 2 * 16384, 8, 3
    sos = po.butter_lowpass_sos(4, 4e6, 250e3)
    ws = [[po.gen_words(n, n0=c * n, seed=31 + s) for s in range(ns)] for c in range(calls)]
    ws[0][4][0] ^= 0x80000000            # first submit, stream 4, chunk 0
    ws[1][1][16384] ^= 0x80000000        # second submit, stream 1, chunk 1
    h_w = [torch.from_numpy(np.stack(ws[c]).view(np.int32).copy()).pin_memory() for c in range(calls)]
    h_o = [torch.zeros((ns, n // D, 2), dtype=torch.float32).pin_memory() for _ in range(calls)]
    with cb.Context(sos=sos, decimation=D, parity=cb.PARITY_F32, n_streams=ns, chunk_bytes=CHUNK) as ctx:
        for c in range(calls):
            ctx.rx_submit(h_w[c], n * 4, h_o[c], stride_bytes=n * 4, filt_stride=n // D, mem=cb.MEM_HOST_PINNED)
        assert ctx.wait()[0] == n
    for s in range(ns):
        ref, _ = _oracle_stream(po, sos, D, [ws[c][s] for c in range(calls)])
        for c in range(calls):
            assert np.all(rel_rms(h_o[c][s].numpy(), ref[c][2]) <= TOL), (s, c)


def test_contexts_on_two_devices_in_one_process(cb, po, torch_cuda):
    """The shared-memory opt-in of the scan kernels is a per-device function attribute: contexts on cuda:0 and
    cuda:1 of one process must both launch (needs a box with >= 2 GPUs)."""
    torch = torch_cuda
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs in one process")
    n, D = 1 << 18, 8
    sos = po.butter_lowpass_sos(4, 4e6, 250e3)
    w = po.gen_words(n)
    ref = po.rx_path_f32(w, sos, D)
    for dev in (0, 1, 0):
        d_w = torch.from_numpy(w.view(np.int32)).to("cuda:%d" % dev)
        out = torch.zeros((n // D, 2), dtype=torch.float32, device="cuda:%d" % dev)
        with cb.Context(device=dev, sos=sos, decimation=D, parity=cb.PARITY_F32) as ctx:
            ctx.rx_submit(d_w, n * 4, out)
            assert ctx.wait()[0] == n
        assert np.all(rel_rms(out.cpu().numpy(), ref) <= TOL), dev
        # full rate: the rerun kernel's attribute as well
        out1 = torch.zeros((n, 2), dtype=torch.float32, device="cuda:%d" % dev)
        with cb.Context(device=dev, sos=sos, decimation=1, parity=cb.PARITY_F32) as ctx:
            ctx.rx_submit(d_w, n * 4, out1)
            assert ctx.wait()[0] == n
        assert np.all(rel_rms(out1.cpu().numpy(), po.rx_path_f32(w, sos, 1)) <= TOL), dev
    torch.cuda.set_device(0)
