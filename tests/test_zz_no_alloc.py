"""GPU test of the allocation-free entry point ``sbx_voronoi_label_noalloc`` (no counterpart in the reference):
called through the C ABI with caller-owned buffers, its labels must equal the
existing entry point's (``sbx_voronoi_plan_execute_range``) bit for bit -- labels are integers, the suite's tolerance
for them is equality -- on calls in a row that share the static scratch array, with different geometries in between so
that stale contents would show.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SBX_ERANGE = -7


def _call(lib_mod, torch, pts, shape, box, mask, planes=None):
    """One call of the new entry point; every buffer is allocated by the caller BEFORE the call."""
    x_lo, x_hi = planes or (0, shape[0])
    labels = torch.full((x_hi - x_lo,) + tuple(shape[1:]), -7, dtype=torch.int32, device="cuda")
    _, pn = lib_mod.i64(shape)
    pb = None
    if box is not None:
        _keep, pb = lib_mod.f64(box)
    m = None if mask is None else mask[x_lo:x_hi].contiguous()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib_mod.lib().sbx_voronoi_label_noalloc(0, len(shape), pn, C.c_void_p(pts.data_ptr()), int(pts.shape[0]), pb,
                                                 C.c_void_p(m.data_ptr()) if m is not None else None,
                                                 C.c_void_p(labels.data_ptr()), x_lo, x_hi, st)
    return rc, labels


def _cases(torch):
    rng = np.random.default_rng(2024)
    out = []
    for shape, S, periodic, masked, planes in [((64, 48, 40), 150, True, True, None), ((128, 96), 300, False, False, None),
                                               ((37, 29, 23), 40, True, False, (5, 30)), ((70, 90), 25, True, True, None)]:
        pts = torch.from_numpy(rng.uniform(-0.1, 1.1, (S, len(shape))) * np.array(shape, dtype=np.float64)).cuda()
        box = tuple(float(n) for n in shape) if periodic else None
        mask = torch.from_numpy((rng.random(shape) > 0.5).astype(np.uint8)).cuda() if masked else None
        out.append((pts, shape, box, mask, planes))
    return out


def test_noalloc_matches_existing_entry_point_on_calls_in_a_row(sbx_lib):
    import torch
    from simubox_b200 import VoronoiCell

    cases = _cases(torch)
    # what the existing entry point gives (plan + sbx_voronoi_plan_execute_range through VoronoiCell.label_grid)
    want = [VoronoiCell.label_grid(p, 0, shape=s, box=b, mask=m, return_tensor=True, planes=pl) for p, s, b, m, pl in cases]
    torch.cuda.synchronize()
    # two passes over geometries that differ in dimension, periodicity and cell layout: every call after the first
    # finds the previous call's points, counters and start table in the shared array
    for rnd in range(2):
        for (p, s, b, m, pl), ref in zip(cases, want):
            rc, got = _call(sbx_lib, torch, p, s, b, m, pl)
            assert rc == 0, sbx_lib.lib().sbx_last_error()
            torch.cuda.synchronize()
            assert torch.equal(got, ref), f"round {rnd}, grid {s}"
    # the same call twice in a row, back to back on the stream without a sync in between
    p, s, b, m, pl = cases[0]
    rc1, a = _call(sbx_lib, torch, p, s, b, m, pl)
    rc2, c = _call(sbx_lib, torch, p, s, b, m, pl)
    torch.cuda.synchronize()
    assert rc1 == 0 and rc2 == 0 and torch.equal(a, want[0]) and torch.equal(c, want[0])


def test_noalloc_baseline_config_and_device_memory_unchanged(sbx_lib):
    """BASELINE config 3 (256^3, 4096 seeds, periodic, uint8 mask: 110592 of the 262144 image points) against the
    whole-grid labelling of the existing path; the free device memory does not shrink across a call."""
    import torch
    from simubox_b200 import VoronoiCell, synth

    shape, box = (256, 256, 256), (256.0, 256.0, 256.0)
    pts = torch.from_numpy(synth.seeds_lattice(shape, 16, 4.0, 1234)).cuda()
    mask = (synth.fields_torch(1, 1, shape, cells=4)[0] > 0.5).to(torch.uint8).contiguous()
    ref = VoronoiCell.label_grid(pts, 0, shape=shape, box=box, mask=mask, return_tensor=True)
    rc, got = _call(sbx_lib, torch, pts, shape, box, mask)  # first call: kernels of this geometry are loaded
    torch.cuda.synchronize()
    assert rc == 0 and torch.equal(got, ref)
    labels = torch.empty_like(ref)
    _, pn = sbx_lib.i64(shape)
    _keep, pb = sbx_lib.f64(box)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    free0, _ = torch.cuda.mem_get_info()
    rc = sbx_lib.lib().sbx_voronoi_label_noalloc(0, 3, pn, C.c_void_p(pts.data_ptr()), 4096, pb, C.c_void_p(mask.data_ptr()),
                                                 C.c_void_p(labels.data_ptr()), 0, 256, st)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert rc == 0 and torch.equal(labels, ref)
    assert free1 >= free0, f"device memory shrank by {free0 - free1} bytes across an allocation-free call"


def test_noalloc_rejects_sizes_beyond_the_static_scratch(sbx_lib):
    import torch

    shape = (64, 64, 64)
    S = 262144 // 27 + 1  # one seed more than the image points of a periodic 3D box allow
    pts = torch.rand((S, 3), dtype=torch.float64, device="cuda") * 64
    rc, _ = _call(sbx_lib, torch, pts, shape, (64.0, 64.0, 64.0), None)
    assert rc == SBX_ERANGE
    msg = sbx_lib.lib().sbx_last_error().decode()
    assert "static scratch" in msg and "sbx_voronoi_plan_execute_range" in msg
    rc, lab = _call(sbx_lib, torch, pts, shape, None, None)  # the same seeds without periodic images fit
    torch.cuda.synchronize()
    assert rc == 0 and int(lab.min()) >= 0
    rc, _ = _call(sbx_lib, torch, pts[:10], shape, None, None, planes=(3, 3))
    assert rc == -1  # SBX_EINVAL: empty plane range
