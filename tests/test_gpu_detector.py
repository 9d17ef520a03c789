"""GPU parity: OrbDetector through the C ABI vs the oracle (restatement, golden vectors, cv2 live)."""
import ctypes
import glob
import os

import numpy as np
import pytest

from oracle import restated as R, synth, cv_ref
from tests.util import compare_orb
from tests.test_oracle_cpu import load_orb, GOLD

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orbx():
    import slam_rs_b200 as s
    return s


def assert_exact(st):
    assert st["missing"] == 0 and st["extra"] == 0, st
    if st["common"]:
        assert st["resp_eq"] == st["common"], st
        assert st["angle_eq"] == st["common"], st
        assert st["pt_eq"] == st["common"] and st["size_eq"] == st["common"], st
        assert st["desc_bits_diff"] == 0, st


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "orb_*.npz"))))
def test_golden_vectors(orbx, path):
    g, ref = load_orb(path)
    h, w, nf = int(g["h"]), int(g["w"]), int(g["nfeatures"])
    det = orbx.OrbDetector(nf, max_width=w, max_height=h)
    kps, desc = det.detect_and_compute(g["image"])
    assert_exact(compare_orb(kps, desc, ref, h, w))
    assert np.all(kps["class_id"] == -1)
    assert np.array_equal(det.level(1), g["level1"])
    assert np.array_equal(det.level(0, blurred=True), g["blur0"])


@pytest.mark.parametrize("h,w,seed,nf,sparse", [
    (376, 1241, 0, 2000, False),      # BASELINE config 0/1 frame
    (376, 1241, 1, 500, False),       # reference-faithful nfeatures (detector.rs:16)
    (1080, 1920, 200, 5000, False),   # BASELINE config 2 frame
    (480, 640, 31, 1000, True),
    (333, 517, 32, 700, False),       # odd sizes: unaligned rows everywhere
])
def test_vs_restated_all_stages(orbx, h, w, seed, nf, sparse):
    img = synth.frame(h, w, seed, sparse)
    det = orbx.OrbDetector(nf, max_width=w, max_height=h)
    kps, desc = det.detect_and_compute(img)
    ref = R.orb_detect_and_compute(img, nf, return_stages=True)
    assert_exact(compare_orb(kps, desc, ref, h, w))
    stg = ref["stages"]
    for l in range(8):
        assert np.array_equal(det.level(l), stg["levels"][l]), "pyramid level %d" % l
        assert np.array_equal(det.level(l, blurred=True), stg["blurred"][l]), "blurred level %d" % l
        x, y, sc = det.candidates(l)
        rx, ry, rs = stg["cand"][l]
        hh, ww = stg["levels"][l].shape
        inb = (rx >= 31) & (rx < ww - 31) & (ry >= 31) & (ry < hh - 31)
        assert set(zip(x.tolist(), y.tolist(), sc.tolist())) == set(
            zip(rx[inb].tolist(), ry[inb].tolist(), rs[inb].tolist())), "FAST/NMS level %d" % l
    # library order: level major, response descending, then (y, x)
    key = np.lexsort((kps["x"], kps["y"], -kps["response"].astype(np.float64), kps["octave"]))
    assert np.array_equal(key, np.arange(len(kps)))
    # detect() == keypoints of detect_and_compute()
    k2 = det.detect(img)
    assert np.array_equal(k2, kps)


@pytest.mark.skipif(not cv_ref.HAVE_CV2, reason="cv2 not importable")
@pytest.mark.parametrize("h,w,seed,nf", [(376, 1241, 7, 2000), (1080, 1920, 201, 5000), (240, 320, 8, 500)])
def test_vs_cv2_live(orbx, h, w, seed, nf):
    img = synth.frame(h, w, seed)
    det = orbx.OrbDetector(nf, max_width=w, max_height=h)
    kps, desc = det.detect_and_compute(img)
    assert_exact(compare_orb(kps, desc, cv_ref.orb_detect_and_compute(img, nf), h, w))


@pytest.mark.skipif(not cv_ref.HAVE_CV2, reason="cv2 not importable")
def test_fast_threshold_param_vs_cv2(orbx):
    import cv2
    img = synth.frame(300, 400, 40)
    det = orbx.OrbDetector(600, max_width=400, max_height=300, fast_threshold=35)
    kps, desc = det.detect_and_compute(img)
    k, d = cv2.ORB_create(nfeatures=600, fastThreshold=35).detectAndCompute(img, None)
    assert_exact(compare_orb(kps, desc, cv_ref._pack(k, d), 300, 400))


def test_reference_invariants(orbx):
    # detector.rs:51-80
    det = orbx.OrbDetector(500, max_width=640, max_height=480, reference_faithful=True)
    assert det.max_features() == 500
    det100 = orbx.OrbDetector(100, max_width=640, max_height=480)
    blank = np.zeros((480, 640), np.uint8)
    assert len(det100.detect(blank)) == 0
    kps, desc = det.detect_and_compute(blank)
    assert len(kps) == 0 and desc.shape == (0, 32)
    # reference_faithful keeps ORB::create_def(): nfeatures stays 500 whatever max_features says
    d2 = orbx.OrbDetector(3000, max_width=640, max_height=480, reference_faithful=True)
    assert d2.max_features() == 3000 and d2.nfeatures == 500
    img = synth.frame(480, 640, 41)
    k, _ = d2.detect_and_compute(img)
    assert len(k) == len(R.orb_detect_and_compute(img, 500)["x"])


def test_small_and_degenerate_images(orbx):
    det = orbx.OrbDetector(100, max_width=640, max_height=480)
    img = synth.frame(70, 200, 3)            # only level 0 can yield keypoints
    kps, desc = det.detect_and_compute(img)
    assert_exact(compare_orb(kps, desc, R.orb_detect_and_compute(img, 100), 70, 200))
    assert np.all(kps["octave"] == 0)
    for shape in [(16, 16), (63, 63), (62, 300)]:
        z = synth.frame(shape[0], shape[1], 4)
        k, d = det.detect_and_compute(z)
        assert_exact(compare_orb(k, d, R.orb_detect_and_compute(z, 100), shape[0], shape[1]))
    # saturated and checkerboard inputs: ties everywhere
    sat = np.full((200, 300), 255, np.uint8)
    assert len(det.detect(sat)) == 0
    yy, xx = np.mgrid[0:240, 0:320]
    chk = (((yy // 16) + (xx // 16)) % 2 * 200 + 20).astype(np.uint8)
    k, d = det.detect_and_compute(chk)
    assert_exact(compare_orb(k, d, R.orb_detect_and_compute(chk, 100), 240, 320))


def dot_lattice(h, w, step):
    """Isolated bright pixels on a flat background: every dot is a FAST corner with the same score and the same Harris
    response, so both retainBest stages keep ALL of them (OpenCV keeps every entry >= the n-th best)."""
    img = np.full((h, w), 30, np.uint8)
    img[4::step, 4::step] = 220
    return img


def test_tie_floods_return_every_tied_keypoint(orbx):
    """cv::ORB returns far more than nfeatures keypoints when a level ties at the selection threshold; so must we
    (detector.rs:26-40 hands back whatever OpenCV returns).  Pinned to cv2 live and to the restatement."""
    import cv2
    det = orbx.OrbDetector(100, max_width=1241, max_height=376, max_batch=3)
    # 704 tied keypoints on level 0 (quota 22): beyond sel_cap, sorted in the spill arena
    img = dot_lattice(240, 320, 8)
    kps, desc = det.detect_and_compute(img)
    ref = R.orb_detect_and_compute(img, 100)
    assert len(kps) == len(ref["x"]) == len(cv2.ORB_create(nfeatures=100).detect(img, None)) == 760
    assert_exact(compare_orb(kps, desc, ref, 240, 320))
    assert_exact(compare_orb(kps, desc, cv_ref.orb_detect_and_compute(img, 100), 240, 320))
    # detect() alone returns the same keypoints
    assert np.array_equal(det.detect(img), kps)
    # ~23 k tied keypoints on level 0: beyond the shared-memory sort (16384) as well
    big = dot_lattice(376, 1241, 4)
    kb, db = det.detect_and_compute(big)
    refb = cv_ref.orb_detect_and_compute(big, 100)
    assert len(kb) == len(refb["x"]) > 20000
    assert_exact(compare_orb(kb, db, refb, 376, 1241))
    # binary noise and block noise (the images a SLAM front end sees on synthetic / thresholded input)
    rng = np.random.default_rng(5)
    for im in ((rng.integers(0, 2, (240, 320)) * 255).astype(np.uint8),
               np.kron(rng.integers(0, 2, (30, 40)), np.ones((8, 8))).astype(np.uint8) * 255):
        k, d = det.detect_and_compute(im)
        assert_exact(compare_orb(k, d, cv_ref.orb_detect_and_compute(im, 100), 240, 320))
    # batch: a flooded frame between two ordinary ones does not disturb its neighbours
    a, c = synth.frame(240, 320, 91), synth.frame(240, 320, 92)
    try:
        det.detect_and_compute_batch(np.stack([a, img, c]), cap=512)
        raise AssertionError("expected ORBX_E_CAPACITY: 760 keypoints do not fit cap=512")
    except orbx.OrbxError as e:
        assert e.code == -2
    out = det.detect_and_compute_batch(np.stack([a, img, c]), cap=1024)
    for (k, d), im in zip(out, (a, img, c)):
        assert_exact(compare_orb(k, d, R.orb_detect_and_compute(im, 100), 240, 320))


def test_async_overflow_is_clamped_and_reported(orbx):
    """The asynchronous (device-pointer) path cannot return a needed size: rows are truncated to cap, n is clamped to
    cap (never an out-of-bounds row count for the consumer) and the next sync reports ORBX_E_CAPACITY once."""
    import torch
    det = orbx.OrbDetector(100, max_width=320, max_height=240, max_batch=2)
    imgs = np.stack([dot_lattice(240, 320, 8), synth.frame(240, 320, 93)])
    ref = det.detect_and_compute_batch(imgs, cap=1024)
    cap = 256
    buf = torch.from_numpy(imgs).cuda()
    kps = torch.zeros((2, cap, 28), dtype=torch.uint8, device="cuda")
    desc = torch.zeros((2, cap, 32), dtype=torch.uint8, device="cuda")
    n = torch.zeros((2,), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    det.detect_and_compute_device(buf, kps, desc, n)
    try:
        det.sync()
        raise AssertionError("expected ORBX_E_CAPACITY from the sync after a truncated asynchronous batch")
    except orbx.OrbxError as e:
        assert e.code == -2 and "760" in str(e)
    nn = n.cpu().numpy()
    assert nn[0] == cap and nn[1] == len(ref[1][0])
    # the rows that fit are the first rows of the full result; the ordinary frame is untouched
    k0 = kps[0].cpu().numpy().view(orbx.KEYPOINT_DTYPE).reshape(-1)
    assert np.array_equal(k0, ref[0][0][:cap]) and np.array_equal(desc[0].cpu().numpy(), ref[0][1][:cap])
    k1 = kps[1, :nn[1]].cpu().numpy().view(orbx.KEYPOINT_DTYPE).reshape(-1)
    assert np.array_equal(k1, ref[1][0])
    det.sync()  # reported once


def test_graph_replay_returns_identical_bytes(orbx):
    """orbx_detector_set_graph: the single-frame host call replayed as one CUDA graph must return the same bytes as the
    stream path, across repeated calls, a change of image size (tables re-uploaded, graphs dropped), detect-only calls
    and a capacity retry."""
    a, b, c = synth.frame(376, 1241, 0), synth.frame(376, 1241, 1), synth.frame(240, 320, 2)
    ref = orbx.OrbDetector(2000, max_width=1241, max_height=376)
    det = orbx.OrbDetector(2000, max_width=1241, max_height=376)
    det.set_graph(True)
    for img in (a, b, a, c, b, c):
        k0, d0 = ref.detect_and_compute(img)
        k1, d1 = det.detect_and_compute(img)
        assert np.array_equal(k0, k1) and np.array_equal(d0, d1) and len(k0) > 0
        assert np.array_equal(ref.detect(img), det.detect(img))
    # more keypoints than the default cap: the graph path reports the needed size like the stream path
    small = orbx.OrbDetector(100, max_width=320, max_height=240)
    small.set_graph(True)
    lat = dot_lattice(240, 320, 8)
    k, d = small.detect_and_compute(lat)
    k2, d2 = orbx.OrbDetector(100, max_width=320, max_height=240).detect_and_compute(lat)
    assert len(k) == 760 and np.array_equal(k, k2) and np.array_equal(d, d2)
    det.set_graph(False)
    k1, d1 = det.detect_and_compute(a)
    assert np.array_equal(k1, ref.detect_and_compute(a)[0])


def test_batch_device_pointers_and_pitch(orbx):
    import torch
    h, w, nf, B = 200, 333, 400, 5
    imgs = np.stack([synth.frame(h, w, 50 + i) for i in range(B)])
    det = orbx.OrbDetector(nf, max_width=w, max_height=h, max_batch=B)
    host = det.detect_and_compute_batch(imgs)
    single = [orbx.OrbDetector(nf, max_width=w, max_height=h).detect_and_compute(im) for im in imgs]
    for (k1, d1), (k2, d2) in zip(host, single):
        assert np.array_equal(k1, k2) and np.array_equal(d1, d2)
    cap = 640
    for pitch in (w, 384):                   # unaligned (odd) pitch and 128-aligned pitch
        buf = torch.zeros((B, h, pitch), dtype=torch.uint8, device="cuda")
        buf[:, :, :w] = torch.from_numpy(imgs).cuda()
        kps = torch.zeros((B, cap, 28), dtype=torch.uint8, device="cuda")
        desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
        n = torch.zeros((B,), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        det.detect_and_compute_device(buf, kps, desc, n, width=w)
        det.sync()
        nn = n.cpu().numpy()
        for i in range(B):
            k = kps[i, :nn[i]].cpu().numpy().view(orbx.KEYPOINT_DTYPE).reshape(-1)
            assert np.array_equal(k, host[i][0]) and np.array_equal(desc[i, :nn[i]].cpu().numpy(), host[i][1])


def test_unaligned_device_frames_are_repacked(orbx):
    """Tightly packed odd-width frames at odd byte offsets inside a device buffer, the last one ending exactly at the end
    of the allocation (csrc/ingest.cu::repack_gray_kernel must not read past it): same bytes out as the host path."""
    import torch
    h, w, nf, B = 97, 203, 300, 3
    imgs = np.stack([synth.frame(h, w, 70 + i) for i in range(B)])
    det = orbx.OrbDetector(nf, max_width=w, max_height=h, max_batch=B)
    host = det.detect_and_compute_batch(imgs)
    cap = 512
    for lead in (0, 1, 3):
        flat = torch.zeros(lead + B * h * w, dtype=torch.uint8, device="cuda")
        flat[lead:] = torch.from_numpy(imgs.reshape(-1)).cuda()
        buf = flat[lead:].view(B, h, w)
        kps = torch.zeros((B, cap, 28), dtype=torch.uint8, device="cuda")
        desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
        n = torch.zeros((B,), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        det.detect_and_compute_device(buf, kps, desc, n, width=w)
        det.sync()
        nn = n.cpu().numpy()
        for i in range(B):
            k = kps[i, :nn[i]].cpu().numpy().view(orbx.KEYPOINT_DTYPE).reshape(-1)
            assert np.array_equal(k, host[i][0]) and np.array_equal(desc[i, :nn[i]].cpu().numpy(), host[i][1])
        assert np.array_equal(det.level(0, frame=B - 1), imgs[B - 1])        # level 0 of the last frame, byte for byte


@pytest.mark.parametrize("w", [204, 208, 332])
def test_tight_aligned_pitch_device_frames_in_place(orbx, w):
    """Device frames whose pitch equals a width that is a multiple of 4 but not of 128 are read in place (no repack,
    no pitch padding behind the rows): same bytes out as the host path."""
    import torch
    h, nf, B = 101, 300, 4
    imgs = np.stack([synth.frame(h, w, 90 + i) for i in range(B)])
    det = orbx.OrbDetector(nf, max_width=w, max_height=h, max_batch=B)
    host = det.detect_and_compute_batch(imgs)
    cap = 512
    buf = torch.from_numpy(imgs).cuda()
    kps = torch.zeros((B, cap, 28), dtype=torch.uint8, device="cuda")
    desc = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
    n = torch.zeros((B,), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    det.detect_and_compute_device(buf, kps, desc, n, width=w)
    det.sync()
    nn = n.cpu().numpy()
    for i in range(B):
        k = kps[i, :nn[i]].cpu().numpy().view(orbx.KEYPOINT_DTYPE).reshape(-1)
        assert np.array_equal(k, host[i][0]) and np.array_equal(desc[i, :nn[i]].cpu().numpy(), host[i][1])


def test_capacity_error_reports_needed_size(orbx):
    img = synth.frame(240, 320, 5)
    det = orbx.OrbDetector(300, max_width=320, max_height=240)
    kps = np.zeros(50, orbx.KEYPOINT_DTYPE)
    desc = np.zeros((50, 32), np.uint8)
    n = ctypes.c_int(0)
    rc = orbx.lib.orbx_detect_and_compute(det._h, img.ctypes.data, 320, 240, 320, kps.ctypes.data, desc.ctypes.data,
                                          50, ctypes.byref(n))
    assert rc == -2 and n.value == 280
    full, fd = det.detect_and_compute(img)
    assert np.array_equal(kps, full[:50]) and np.array_equal(desc, fd[:50])
    # too large for the handle
    big = np.zeros((300, 400), np.uint8)
    with pytest.raises(orbx.OrbxError):
        det.detect(big)


def test_determinism(orbx):
    img = synth.frame(376, 1241, 60)
    det = orbx.OrbDetector(2000, max_width=1241, max_height=376)
    a = det.detect_and_compute(img)
    for _ in range(3):
        b = det.detect_and_compute(img)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_bgr_ingest(orbx):
    """Frame ingest in front of S1 (visual_odometry.rs:93-99): 8UC3 BGR input is gray-converted on the device with
    OpenCV's 15-bit fixed-point weights; level 0 and all results must equal the run on the oracle's gray image."""
    import torch
    from tests.test_oracle_cpu import bgr_frame
    for h, w, nf in ((200, 333, 400), (240, 320, 300)):            # odd width: unaligned 3*w rows
        bgr = bgr_frame(h, w, 60)
        gray = R.bgr_to_gray(bgr)
        det = orbx.OrbDetector(nf, max_width=w, max_height=h, max_batch=3)
        k1, d1 = det.detect_and_compute(bgr)
        assert np.array_equal(det.level(0), gray)
        k2, d2 = det.detect_and_compute(gray)
        assert np.array_equal(k1, k2) and np.array_equal(d1, d2) and len(k1) > 50
        assert np.array_equal(det.detect(bgr), k2)
        batch = np.stack([bgr, bgr[::-1].copy(), bgr[:, ::-1].copy()])
        host = det.detect_and_compute_batch_bgr(batch)
        for i in range(3):
            kg, dg = det.detect_and_compute(R.bgr_to_gray(batch[i]))
            assert np.array_equal(host[i][0], kg) and np.array_equal(host[i][1], dg)
        cap = 640
        dev = torch.from_numpy(batch).cuda()
        kps = torch.zeros((3, cap, 28), dtype=torch.uint8, device="cuda")
        desc = torch.zeros((3, cap, 32), dtype=torch.uint8, device="cuda")
        n = torch.zeros((3,), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        det.detect_and_compute_device_bgr(dev, kps, desc, n)
        det.sync()
        nn = n.cpu().numpy()
        for i in range(3):
            k = kps[i, :nn[i]].cpu().numpy().view(orbx.KEYPOINT_DTYPE).reshape(-1)
            assert np.array_equal(k, host[i][0]) and np.array_equal(desc[i, :nn[i]].cpu().numpy(), host[i][1])
