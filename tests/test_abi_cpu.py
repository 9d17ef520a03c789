"""CPU: the C-ABI library loads and exports every symbol include/orbx.h declares; the host mirror fails
loudly without a GPU (no CPU fallback); multi-rank sharding logic under gloo (world_size 2)."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "orbx.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_header_symbol():
    import slam_rs_b200 as s
    lib = ctypes.CDLL(s.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 30
    for name in syms:
        assert hasattr(lib, name), "liborbx.so does not export %s" % name
    # and the ctypes binding covers the same set
    assert set(syms) == set(s.SYMBOLS), set(syms) ^ set(s.SYMBOLS)


def test_rust_ffi_crate_mirrors_the_header():
    """rust/orbx-sys/src/lib.rs is GENERATED from include/orbx.h (scripts/gen_rust_sys.py): it must be current, declare
    exactly the header's symbols, and its build.rs must compile every .cu the Makefile links (the crate
    omitted sequence.cu and 20 symbols)."""
    import re
    import subprocess
    import sys
    rc = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "gen_rust_sys.py"), "--check"]).returncode
    assert rc == 0, "rust/orbx-sys/src/lib.rs is stale: run python scripts/gen_rust_sys.py"
    rs = open(os.path.join(ROOT, "rust", "orbx-sys", "src", "lib.rs")).read()
    assert set(re.findall(r"pub fn (orbx_\w+)", rs)) == set(header_symbols())
    mk = open(os.path.join(ROOT, "slam-rs_b200", "csrc", "Makefile")).read()
    srcs = set(re.search(r"^SRCS\s*:=\s*(.*)$", mk, re.M).group(1).split())
    on_disk = {f for f in os.listdir(os.path.join(ROOT, "slam-rs_b200", "csrc")) if f.endswith(".cu")}
    assert srcs == on_disk, srcs ^ on_disk
    build_rs = open(os.path.join(ROOT, "rust", "orbx-sys", "build.rs")).read()
    assert "Makefile" in build_rs and "SRCS" in build_rs          # the crate takes its source list from the Makefile
    safe = open(os.path.join(ROOT, "rust", "orbx", "src", "lib.rs")).read()
    for name in ("fn knn_match", "fn ratio_test", "fn new_cross_check", "pub struct Map", "fn find_matches",
                 "pub struct ShardedMatcher", "fn new_reference_faithful"):
        assert name in safe, name
    types = set(re.findall(r"pub (?:enum|struct) (orbx_\w+)", rs))
    for call in set(re.findall(r"sys::(orbx_\w+)", safe)) - types:   # every call the safe crate makes exists in the header
        assert call in header_symbols(), call


def test_struct_layouts_match_opencv():
    import slam_rs_b200 as s
    from slam_rs_b200 import _ffi
    assert ctypes.sizeof(_ffi.Keypoint) == 28 and s.KEYPOINT_DTYPE.itemsize == 28
    assert ctypes.sizeof(_ffi.DMatch) == 16 and s.DMATCH_DTYPE.itemsize == 16
    p = _ffi.Params()
    s.lib.orbx_default_params(ctypes.byref(p))
    # ORB::create_def() (detector.rs:16)
    assert (p.nfeatures, p.nlevels, p.edge_threshold, p.first_level, p.wta_k, p.score_type, p.patch_size,
            p.fast_threshold) == (500, 8, 31, 0, 2, 0, 31, 20)
    assert abs(p.scale_factor - 1.2) < 1e-6


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import slam_rs_b200 as s
    with pytest.raises(s.OrbxError) as e:
        s.OrbDetector(500)
    assert e.value.code == -3 and "no CPU fallback" in str(e.value)
    with pytest.raises(s.OrbxError):
        s.FeatureMatcher()
    cam = s.CameraIntrinsics.kitti()
    for make in (lambda: s.Map(cam, capacity=16), lambda: s.Triangulator(cam, max_points=16)):
        with pytest.raises(s.OrbxError) as e:
            make()
        assert e.value.code == -3
    from slam_rs_b200.frontend import NativeSequence
    with pytest.raises(s.OrbxError) as e:
        NativeSequence(640, 480, max_frames=4)
    assert e.value.code == -3
    h = ctypes.c_void_p()                      # (the PoseEstimator mirror creates a FeatureMatcher first: go to the ABI)
    assert s.lib.orbx_pose_estimator_create(0, 64, 1, cam.fx, cam.fy, cam.cx, cam.cy, ctypes.byref(h)) == -3
    assert b"no CPU fallback" in s.lib.orbx_last_error()


def test_bad_arguments_are_errors_not_crashes():
    import slam_rs_b200 as s
    h = ctypes.c_void_p()
    assert s.lib.orbx_detector_create(None, 0, 4, 4, 1, ctypes.byref(h)) == -1      # size out of range
    from slam_rs_b200 import _ffi
    p = _ffi.Params()
    s.lib.orbx_default_params(ctypes.byref(p))
    p.nlevels = 4
    assert s.lib.orbx_detector_create(ctypes.byref(p), 0, 640, 480, 1, ctypes.byref(h)) == -1
    assert b"create_def" in s.lib.orbx_last_error()
    assert s.lib.orbx_detector_max_features(None) == -1


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "slam-rs_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inc", ".h")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "import oracle" not in txt and "from oracle" not in txt and "import cv2" not in txt, f


_WORKER = r'''
import os, sys
sys.path.insert(0, %(root)r)
import numpy as np, torch, torch.distributed as dist
from oracle import restated as R, synth
from slam_rs_b200 import sharding
from slam_rs_b200.mapmatch import MapToFrameMatcher
from slam_rs_b200.feature import DMATCH_DTYPE
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")


class OracleMatcher:
    # stand-in for the per-rank CUDA matcher (no GPU here): same match_device / filter_good_matches contract,
    # query_idx local to the slice it is given - exactly what the kernel returns
    def match_device(self, q, nq, t, nt, out, n_out=None, k=1):
        idx, d, c = R.knn(q.numpy()[:nq], t.numpy()[:nt], k)
        rec = np.zeros((nq, k), DMATCH_DTYPE)
        rec["query_idx"] = np.arange(nq)[:, None]
        rec["train_idx"], rec["distance"] = idx, d
        out[:nq] = torch.from_numpy(rec.view(np.uint8).reshape((nq, k, 16) if k == 2 else (nq, 16)))
        if n_out is not None and k == 2:
            n_out[:nq] = torch.from_numpy(c.astype(np.int32))

    def filter_good_matches(self, m, ratio):
        keep = R.filter_good_matches(np.stack([m["query_idx"], m["train_idx"], m["distance"]], 1), ratio)
        return keep


m, f = synth.map_and_frame_descriptors(4000, 301, f=0)
# the PRODUCT class (slicing, query_idx offset, padded all-gather) over a real world-size-2 group
mm = MapToFrameMatcher(device="cpu", rank=rank, world=world, dist=dist, matcher=OracleMatcher())
mm.set_train(m.copy() if rank == 0 else None, rows=len(m))
q = torch.from_numpy(f.copy()) if rank == 0 else None
rec, cnt = mm.knn2(q, nq=len(f))
rec = rec.numpy().view(DMATCH_DTYPE).reshape(len(f), 2)
ri, rd, rc = R.knn(f, m, 2)
assert np.array_equal(rec["query_idx"], np.repeat(np.arange(len(f))[:, None], 2, 1))   # global rows on every rank
assert np.array_equal(rec["train_idx"], ri) and np.array_equal(rec["distance"], rd.astype(np.float32))
assert np.array_equal(cnt.numpy(), rc)
# reference orientation (map.rs:107-108): the resident rows are the queries, the frame is broadcast
fr = torch.from_numpy(f[:97].copy()) if rank == 0 else torch.zeros((97, 32), dtype=torch.uint8)
good = mm.match_train_as_query(fr, 2.0)
if rank == 0:
    want = R.filter_good_matches(R.match_k1(m, f[:97]), 2.0)
    assert np.array_equal(np.asarray(good)[:, :2].astype(np.int64), np.asarray(want)[:, :2].astype(np.int64))
# an empty frame is an empty result on every rank, with no collective left half-issued (ADVICE r1)
empty = mm.match_train_as_query(torch.zeros((0, 32), dtype=torch.uint8), 2.0)
assert (rank == 0 and len(empty) == 0) or (rank != 0 and empty is None)
# frame sharding for extraction
fr = sharding.frame_shard(10, rank, world)
cnt = torch.tensor([len(fr)]); dist.all_reduce(cnt)
assert int(cnt) == 10
dist.barrier()
if rank == 0:
    print("OK")
dist.destroy_process_group()
'''


def test_query_sharding_world2_gloo(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER % {"root": ROOT})
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29731", WORLD_SIZE="2")
    procs = []
    for r in range(2):
        e = dict(env, RANK=str(r), LOCAL_RANK=str(r))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=e, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=300)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "OK" in outs[0]


def test_bench_reference_arm_prints_exactly_one_json_line():
    """bench.py --impl reference (the reference's OpenCV path on the host cores): stdout is ONE JSON line with the
    contract's keys; nothing else may reach stdout (the NCCL banner once did)."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--ref-frames", "3"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "frames/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["n_gpus"] == 1 and d["steps"] == 1 and d["vs_baseline"] is None and d["dtype"] == "u8"
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert "workload" in d["config"]
