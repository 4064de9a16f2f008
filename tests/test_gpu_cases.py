"""GPU: adversarial inputs, reference-golden digests, capacity/err behaviour, device-resident inputs."""
import json
import os

import numpy as np
import pytest

import cases
from hls_final_visual_odometry_b200 import Engine, VoError
from hls_final_visual_odometry_b200 import api
from hls_final_visual_odometry_b200.synth import make_sequence
from oracle.oracle_lib import Oracle, features_digest, matches_digest
from test_gpu_parity import assert_features_equal, assert_matches_equal

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = json.load(open(os.path.join(HERE, "golden", "golden.json")))


# 1241x376 (V = 9) and 2048x1024 (V = 23: K2 runs in row groups of 8 with chunked windows and the one-column fallback on
# `dense`) are the BASELINE sizes; 1024x284 is the reference's own geometry
@pytest.mark.parametrize("size", [(1024, 284), (500, 150), (1241, 376), (2048, 1024)], ids=lambda s: f"{s[0]}x{s[1]}")
@pytest.mark.parametrize("name", sorted(cases.CASES))
def test_adversarial_vs_oracle(name, size):
    w, h = size
    prev, curr = cases.CASES[name](w, h, 0)
    orc = Oracle(w, h)
    eng = Engine(w, h, max_frames=2, max_pairs=1)
    try:
        fp = orc.compute_features(prev)[:2]
        fc = orc.compute_features(curr)[:2]
        eng.batch_extract_host(0, np.stack([prev, curr]))
        assert_features_equal(eng.slot_dump_features(0), fp, f"{name} prev")
        assert_features_equal(eng.slot_dump_features(1), fc, f"{name} curr")
        eng.batch_match([0], [1], max_features=2, bw=50.0, bh=50.0)
        want, _ = orc.matching(fp[0], fp[1], fc[0], fc[1], cap=1 << 18)
        assert_matches_equal(eng.batch_get_matches(0, 0), want, name)
        assert_matches_equal(eng.batch_get_matches(0, 1), orc.bucket_features(want, 2, 50.0, 50.0), name + " bucketed")
    finally:
        eng.close()


# live bin size: the tile kernel's other instantiations, incl. the cases the reference leaves undefined
@pytest.mark.parametrize("geo", [(1241, 376, 45, 45), (1024, 284, 60, 30), (1241, 376, 90, 30), (2048, 1024, 45, 45)],
                         ids=lambda g: f"{g[0]}x{g[1]}_b{g[2]}x{g[3]}")
@pytest.mark.parametrize("name", ["holes", "dead_zone_only", "low_contrast", "dense", "periodic"])
def test_bin_sizes_vs_oracle(name, geo):
    w, h, bw, bh = geo
    prev, curr = cases.CASES[name](w, h, 0)
    orc = Oracle(w, h, bw, bh)
    eng = Engine(w, h, max_frames=2, max_pairs=1, bin_w=bw, bin_h=bh)
    try:
        assert (eng.U, eng.V, eng.CAP) == (orc.U, orc.V, orc.CAP)
        fp = orc.compute_features(prev)[:2]
        fc = orc.compute_features(curr)[:2]
        eng.batch_extract_host(0, np.stack([prev, curr]))
        assert_features_equal(eng.slot_dump_features(0), fp, f"{name} prev")
        assert_features_equal(eng.slot_dump_features(1), fc, f"{name} curr")
        eng.batch_match([0], [1], max_features=2, bw=50.0, bh=50.0)
        want, _ = orc.matching(fp[0], fp[1], fc[0], fc[1], cap=1 << 18)
        assert_matches_equal(eng.batch_get_matches(0, 0), want, name)
    finally:
        eng.close()


# live search radius: SEARCH_RAD_U x SEARCH_RAD_V other than three bins, incl. the cases the reference leaves
# undefined, a radius of a few pixels and one beyond the three-bin range (which then limits the search, src/myMatch.cpp:90-99)
@pytest.mark.parametrize("geo", [(1241, 376, 96, 64), (1024, 284, 40, 150), (2048, 1024, 30, 200), (640, 200, 300, 300), (1241, 376, 7, 5)],
                         ids=lambda g: f"{g[0]}x{g[1]}_r{g[2]}x{g[3]}")
@pytest.mark.parametrize("name", ["holes", "low_contrast", "dense", "periodic", "big_motion"])
def test_search_radius_vs_oracle(name, geo):
    w, h, ru, rv = geo
    prev, curr = cases.CASES[name](w, h, 0)
    orc = Oracle(w, h, 60, 45, ru, rv)
    eng = Engine(w, h, max_frames=2, max_pairs=1, match_radius_u=ru, match_radius_v=rv)
    try:
        fp = orc.compute_features(prev)[:2]
        fc = orc.compute_features(curr)[:2]
        eng.batch_extract_host(0, np.stack([prev, curr]))
        eng.batch_match([0], [1], max_features=2, bw=50.0, bh=50.0)
        want, _ = orc.matching(fp[0], fp[1], fc[0], fc[1], cap=1 << 18)
        assert_matches_equal(eng.batch_get_matches(0, 0), want, name)
        assert_matches_equal(eng.batch_get_matches(0, 1), orc.bucket_features(want, 2, 50.0, 50.0), f"{name} bucketed")
    finally:
        eng.close()


def test_search_radius_is_validated():
    from hls_final_visual_odometry_b200.api import VoError
    with pytest.raises(VoError):
        Engine(640, 200, match_radius_u=5000)
    with pytest.raises(VoError):
        Engine(640, 200, match_radius_v=-1)


def test_one_sequence_split_into_frame_ranges_equals_the_unsplit_run():
    """SURVEY.md section 8 row E, second half: contiguous frame ranges of ONE sequence with a one-frame overlap, each range on
    its own engine (one per rank in tools/split_sequence.py; here three engines in one process) -- every pair's raw and
    bucketed list is bit-identical to the single-engine run."""
    import importlib.util
    import os
    from hls_final_visual_odometry_b200 import sharding
    spec = importlib.util.spec_from_file_location("split_sequence", os.path.join(os.path.dirname(__file__), "..", "tools", "split_sequence.py"))
    ss = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ss)
    w, h, n = 640, 200, 23
    seq = make_sequence(w, h, n, seed=2)
    whole, _ = ss.run_range(seq, 0, n, batch=8)
    assert sorted(whole) == list(range(1, n)) and all(v[0] > 100 for v in whole.values())
    for world in (2, 3):
        merged = {}
        for r in range(world):
            f0, f1 = sharding.split_frame_ranges(n, world, r)
            part, _ = ss.run_range(seq, f0, f1, batch=8)
            assert not (set(part) & set(merged))
            merged.update(part)
        assert merged == whole


def test_engines_on_two_devices_in_one_process():
    """the > 48 KB shared-memory opt-in of K1/K2/K4 is a per-device attribute: an engine on a second GPU of the same
    process must work like the first (round-1 bug: a process-wide `static bool attr_set`)"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    w, h = 1241, 376
    seq = make_sequence(w, h, 2, seed=3)
    orc = Oracle(w, h)
    fp = orc.compute_features(seq[0])[:2]
    fc = orc.compute_features(seq[1])[:2]
    want, _ = orc.matching(fp[0], fp[1], fc[0], fc[1], cap=1 << 18)
    engines = [Engine(w, h, max_frames=2, max_pairs=1, device=d) for d in (0, 1)]
    try:
        for eng in engines:                       # interleaved: both engines alive, both devices used in turn
            eng.batch_extract_host(0, seq)
        for d, eng in enumerate(engines):
            eng.batch_match([0], [1], max_features=2, bw=50.0, bh=50.0)
            assert_features_equal(eng.slot_dump_features(1), fc, f"device {d}")
            assert_matches_equal(eng.batch_get_matches(0, 0), want, f"device {d}")
            assert_matches_equal(eng.batch_get_matches(0, 1), orc.bucket_features(want, 2, 50.0, 50.0), f"device {d} bucketed")
    finally:
        for eng in engines:
            eng.close()


GOLD_IDS = [(tag, name) for tag, g in GOLD["geometries"].items() if tag != "native" for name in g["cases"]]


@pytest.mark.parametrize("tag,name", GOLD_IDS, ids=[f"{t}-{n}" for t, n in GOLD_IDS])
def test_cuda_matches_reference_golden(tag, name):
    """the CUDA path against digests the REFERENCE produced (no oracle in between)"""
    from test_oracle_golden import _inputs
    geom = GOLD["geometries"][tag]
    rec = geom["cases"][name]
    prev, curr = _inputs(geom, name)
    eng = Engine(geom["W"], geom["H"], max_frames=2, max_pairs=1, bin_w=geom.get("bin_w", 0), bin_h=geom.get("bin_h", 0),
                 match_radius_u=geom.get("rad_u", 0), match_radius_v=geom.get("rad_v", 0))
    try:
        eng.batch_extract_host(0, np.stack([prev, curr]))
        assert features_digest(*eng.slot_dump_features(0)) == rec["features_prev"]
        assert features_digest(*eng.slot_dump_features(1)) == rec["features_curr"]
        eng.batch_match([0], [1], max_features=2, bw=50.0, bh=50.0)
        raw = eng.batch_get_matches(0, 0)
        assert len(raw) == rec["n_raw"] and matches_digest(raw) == rec["raw_matches"]
        if "bucketed_raw_2_50_50" in rec and rec.get("n_after_outliers") == rec["n_raw"]:
            assert matches_digest(eng.batch_get_matches(0, 1)) == rec["bucketed_2_50_50"]
    finally:
        eng.close()


def test_reference_array_stride_160():
    """the reference's own MAX_FP_IN_BIN = 160 layout (src/size.h:18) round-trips"""
    w, h = 1024, 284
    seq = make_sequence(w, h, 2, seed=0)
    orc = Oracle(w, h)
    eng = Engine(w, h)
    try:
        m, n = eng.compute_features(seq[0], cap_stride=160)
        wm, wn, _ = orc.compute_features(seq[0])
        assert m.shape[3] == 160
        assert_features_equal((m, n), (wm, wn), "stride 160")
        m1, n1 = eng.compute_features(seq[1], cap_stride=160)
        got = eng.matching(m, n, m1, n1)
        wm1, wn1, _ = orc.compute_features(seq[1])
        want, _ = orc.matching(wm, wn, wm1, wn1)
        assert_matches_equal(got, want, "stride 160 matching")
    finally:
        eng.close()


def test_capacity_errors_are_reported():
    w, h = 500, 150
    seq = make_sequence(w, h, 2, seed=0)
    eng = Engine(w, h, max_matches=100)
    try:
        with pytest.raises(VoError) as ei:
            eng.compute_features(seq[0], cap_stride=8)   # lists are longer than 8: the reference would overflow
        assert ei.value.status == api.VO_ERR_CAPACITY
        eng.push_back(seq[0]); eng.push_back(seq[1])
        with pytest.raises(VoError) as ei:
            eng.match_features(0)                        # 2324 raw matches do not fit max_matches=100
        assert ei.value.status == api.VO_ERR_CAPACITY
    finally:
        eng.close()


def test_push_back_error_behaviour():
    eng = Engine(500, 150)
    try:
        img = np.zeros((150, 500), np.uint8)
        import ctypes as C
        dims = (C.c_int32 * 3)(500, 150, 400)  # bytes per line < width: src/matcher.cpp:59-62
        st = eng.lib.vo_push_back(eng.h, img.ctypes.data_as(C.c_void_p), None, dims, 0)
        assert st == api.VO_ERR_INVALID and b"Image dimension mismatch" in eng.lib.vo_last_error()
        dims = (C.c_int32 * 3)(500, 150, 500)
        assert eng.lib.vo_push_back(eng.h, None, None, dims, 0) == api.VO_ERR_INVALID
        dims = (C.c_int32 * 3)(640, 150, 640)
        assert eng.lib.vo_push_back(eng.h, img.ctypes.data_as(C.c_void_p), None, dims, 0) == api.VO_ERR_INVALID
        # replace=True overwrites the current frame only
        seq = make_sequence(500, 150, 3, seed=0)
        orc = Oracle(500, 150)
        eng.push_back(seq[0]); eng.push_back(seq[1]); eng.push_back(seq[2], replace=True)
        assert_features_equal(eng.dump_features(0), orc.compute_features(seq[0])[:2], "prev after replace")
        assert_features_equal(eng.dump_features(1), orc.compute_features(seq[2])[:2], "curr after replace")
    finally:
        eng.close()


@pytest.mark.parametrize("pitch_pad", [0, 7, 12], ids=["tight", "odd-pitch", "tma-aligned"])
def test_device_resident_input(pitch_pad):
    """HBM-resident frames: read in place through a caller tensor map when 16-byte aligned,
    otherwise staged device-to-device; both give identical features"""
    import torch
    w, h = 500, 150
    pitch = w + pitch_pad                     # 500, 507, 512
    seq = make_sequence(w, h, 3, seed=4)
    buf = np.zeros((3, h, pitch), np.uint8)
    buf[:, :, :w] = seq
    d = torch.from_numpy(buf).cuda()
    orc = Oracle(w, h)
    eng = Engine(w, h, max_frames=3, max_pairs=2)
    try:
        eng.batch_extract_device(0, d.data_ptr(), 3, h * pitch, pitch)
        for f in range(3):
            assert_features_equal(eng.slot_dump_features(f), orc.compute_features(seq[f])[:2], f"slot {f}")
    finally:
        eng.close()


def test_full_size_properties():
    """BASELINE sizes through size-independent properties (2048x1024 and 1241x376)"""
    for w, h in [(2048, 1024), (1241, 376)]:
        seq = make_sequence(w, h, 2, seed=9)
        eng = Engine(w, h, max_frames=2, max_pairs=1)
        try:
            eng.batch_extract_host(0, seq)
            eng.batch_match([0], [1], max_features=2, bw=50.0, bh=50.0)
            raw = eng.batch_get_matches(0, 0)
            b = eng.batch_get_matches(0, 1)
            assert len(raw) > 10000
            cur = raw["u1c"].astype(np.int64) * 65536 + raw["v1c"].astype(np.int64)
            assert len(np.unique(cur)) == len(cur)
            assert (np.abs(raw["u1c"] - raw["u1p"]) <= 180).all() and (np.abs(raw["v1c"] - raw["v1p"]) <= 135).all()
            assert (raw["u1c"] < 60 * (eng.U - 1)).all()
            flow = np.stack([raw["u1c"] - raw["u1p"], raw["v1c"] - raw["v1p"]], 1)
            assert (flow == [-3, -1]).all(axis=1).mean() > 0.99
            cols = int(raw["u1c"].max() // 50) + 1
            ids = (b["v1c"] // 50).astype(int) * cols + (b["u1c"] // 50).astype(int)
            assert (np.diff(ids) >= 0).all() and np.bincount(ids).max() <= 2
            # swapping the roles of the two frames reverses every match (circular consistency)
            eng.batch_match([1], [0], max_features=0)
            rev = eng.batch_get_matches(0, 0)
            fwd = set(zip(raw["u1p"].tolist(), raw["v1p"].tolist(), raw["u1c"].tolist(), raw["v1c"].tolist()))
            bwd = set(zip(rev["u1c"].tolist(), rev["v1c"].tolist(), rev["u1p"].tolist(), rev["v1p"].tolist()))
            assert len(fwd & bwd) > 0.98 * len(fwd)
        finally:
            eng.close()


def test_cpp_shim_demo(tmp_path):
    """the C++ drop-in (Matcher / VisualOdometryMono::process, reference class surface) driven by the
    reference's demo loop over raw I1_%06d.dat frames"""
    import subprocess
    from hls_final_visual_odometry_b200 import build
    build.build_all()
    w, h, n = 500, 150, 4
    seq = make_sequence(w, h, n, seed=6)
    for f in range(n):
        seq[f].tofile(str(tmp_path / f"I1_{f:06d}.dat"))
    out = subprocess.run([build.DEMO, str(tmp_path), str(w), str(h), str(n)], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    import re
    counts = [int(re.search(r"Matches: (\d+)", l).group(1)) for l in out.stdout.splitlines() if "Matches:" in l]
    # process() = pushBack -> matchFeatures (GPU matching + host removeOutliers) -> bucketFeatures(2, 50, 50):
    # src/viso_mono.cpp:33-39; expected counts from the oracle's matching + bucketing around the host filter
    from hls_final_visual_odometry_b200.api import remove_outliers_host
    orc = Oracle(w, h)
    feats = [orc.compute_features(seq[f])[:2] for f in range(n)]
    want = [0]
    for f in range(1, n):
        raw, _ = orc.matching(feats[f - 1][0], feats[f - 1][1], feats[f][0], feats[f][1])
        want.append(len(orc.bucket_features(remove_outliers_host(raw), 2, 50.0, 50.0)))
    assert counts == want
    assert "mean flow: -3 -1" in out.stdout
    # I/O adapter of SURVEY.md row F4: the same frames as KITTI-style PNG files (image_0/%06d.png, mixed row filters), size
    # taken from the files -- the demo must print exactly what it printed for the raw frames
    from hls_final_visual_odometry_b200 import image_io
    kitti = tmp_path / "kitti" / "image_0"
    kitti.mkdir(parents=True)
    rng = np.random.default_rng(0)
    for f in range(n):
        (kitti / f"{f:06d}.png").write_bytes(image_io.encode_png(seq[f], filters=rng.integers(0, 5, h)))
    out_png = subprocess.run([build.DEMO, str(tmp_path / "kitti"), "0", "0", str(n)], capture_output=True, text=True, timeout=120)
    assert out_png.returncode == 0, out_png.stderr
    assert out_png.stdout == out.stdout


def test_matcher_facade_matches_reference_golden():
    """pushBack -> matchFeatures (GPU matching + host removeOutliers) -> bucketFeatures -> getMatches over 4
    frames, against digests the reference's own Matcher produced (golden 'matcher_flow')."""
    flow = GOLD["matcher_flow"]
    assert flow["geometry"] == "native"
    W, H = 1024, 284
    seq = make_sequence(W, H, 4, seed=flow["seed"])
    m = api.Matcher(W, H)
    try:
        for f, rec in enumerate(flow["frames"]):
            m.pushBack(seq[f], [W, H, W])
            m.matchFeatures(0)
            after = m.getMatches()
            assert len(after) == rec["n_after_outliers"], f"frame {f}"
            assert matches_digest(after) == rec["after_outliers"], f"frame {f}"
            m.bucketFeatures(2, 50.0, 50.0)
            b = m.getMatches()
            assert len(b) == rec["n_bucketed"] and matches_digest(b) == rec["bucketed"], f"frame {f} bucketed"
    finally:
        m.engine.close()


def test_engine_remove_outliers_with_real_outliers():
    """vo_set_matches -> vo_remove_outliers -> vo_bucket_features on a list with gross outliers, against
    the reference-made golden of the host stage and the oracle's bucketing of that result"""
    lst = cases.outlier_lists(api.P_MATCH_DTYPE, 0)["kitti_10k"]
    rec = GOLD["outliers"]["kitti_10k_seed0"]
    eng = Engine(1241, 376)
    try:
        eng.set_matches(lst)
        eng.remove_outliers()
        kept = eng.get_matches()
        assert len(kept) == rec["n_out"] and matches_digest(kept) == rec["out"]
        eng.bucket_features(2, 50.0, 50.0)
        want = Oracle(1241, 376).bucket_features(kept, 2, 50.0, 50.0)
        assert_matches_equal(eng.get_matches(), want, "bucketed after removeOutliers")
    finally:
        eng.close()


def test_back_to_back_async_batches_keep_their_own_pair_tables():
    """A pipelined caller (bench.py's e2e loop) enqueues the next batch while the previous one has not started: the pinned
    pair/edge tables behind the asynchronous copies must not be shared between calls in flight."""
    from hls_final_visual_odometry_b200 import pinned_empty
    w, h, n = 500, 150, 4
    seq = make_sequence(w, h, n, seed=12)
    seq[3] = seq[3][::-1].copy()                             # an unrelated frame: pairs (0,1) and (2,3) give different counts
    eng = Engine(w, h, max_frames=n, max_pairs=1)
    try:
        eng.batch_extract_host(0, seq)
        want = {}
        for pr in ((0, 1), (2, 3)):
            eng.batch_match([pr[0]], [pr[1]])
            want[pr] = int(eng.batch_counts(1)[0][0])
        assert want[(0, 1)] != want[(2, 3)]
        counts = pinned_empty((16, 2), np.int32)
        counts[:] = -1
        eng.synchronize()
        for it in range(16):                                 # no host synchronisation between the calls
            pr = (0, 1) if it % 2 == 0 else (2, 3)
            eng.batch_match([pr[0]], [pr[1]])
            eng.batch_download_async(1, 0, counts[it], None, 0, 0)
        eng.synchronize()
        assert [int(c) for c in counts[:, 0]] == [want[(0, 1)] if it % 2 == 0 else want[(2, 3)] for it in range(16)]
    finally:
        eng.close()


def _noisy(seq, amp, seed):
    rng = np.random.default_rng(seed)
    return np.clip(seq.astype(np.int16) + rng.integers(-amp, amp + 1, seq.shape, dtype=np.int16), 0, 255).astype(np.uint8)


@pytest.mark.parametrize("amp", [1, 3, 8, 24])
@pytest.mark.parametrize("size", [(1241, 376), (640, 200)], ids=lambda s: f"{s[0]}x{s[1]}")
def test_matching_under_sensor_noise_vs_oracle(size, amp):
    """Frames that are NOT bit-identical translations: no origin has a zero-SAD partner, the warps of K2 filter with
    both signature words and keep poorly seeded origins on the lane path (csrc/vo_match.cu).  Bit-exact vs the oracle."""
    w, h = size
    seq = _noisy(make_sequence(w, h, 3, seed=21), amp, 1000 + amp)
    orc = Oracle(w, h)
    eng = Engine(w, h, max_frames=3, max_pairs=2)
    try:
        eng.batch_extract_host(0, seq)
        eng.batch_match([0, 1], [1, 2], max_features=2, bw=50.0, bh=50.0)
        f = [orc.compute_features(seq[i])[:2] for i in range(3)]
        for i in range(2):
            want, _ = orc.matching(*f[i], *f[i + 1], cap=1 << 18)
            assert_matches_equal(eng.batch_get_matches(i, 0), want, f"noise {amp} pair {i}")
            assert_matches_equal(eng.batch_get_matches(i, 1), orc.bucket_features(want, 2, 50.0, 50.0), f"noise {amp} pair {i} bucketed")
    finally:
        eng.close()


@pytest.mark.parametrize("size", [(1241, 376), (2048, 1024)], ids=lambda s: f"{s[0]}x{s[1]}")
def test_pruned_equals_exhaustive_on_mixed_noise_batch(size):
    """One batch whose pairs carry different noise levels (0, 1, 2, 5, 12, 40 grey levels): one- and two-word warps, seed
    handovers, budget handovers and chunked windows in the same launch; pruned search == every-SAD search."""
    w, h = size
    base = make_sequence(w, h, 7, seed=33)
    amps = [0, 0, 1, 2, 5, 12, 40]
    seq = np.stack([_noisy(base[i:i + 1], a, 50 + i)[0] if a else base[i] for i, a in enumerate(amps)])
    prev = np.arange(6, dtype=np.int32)
    got = []
    for exhaustive in (False, True):
        eng = Engine(w, h, max_frames=7, max_pairs=6, match_exhaustive=exhaustive)
        try:
            eng.batch_extract_host(0, seq)
            eng.batch_match(prev, prev + 1, max_features=0)
            got.append([eng.batch_get_matches(i, 0) for i in range(6)])
        finally:
            eng.close()
    for i in range(6):
        assert len(got[0][i]) > 100, f"pair {i}"
        assert matches_digest(got[0][i]) == matches_digest(got[1][i]), f"pair {i} (noise {amps[i]} -> {amps[i + 1]})"


def test_random_stress_pruned_equals_exhaustive():
    """tools/stress_match.py: random sizes, flows up to the window limits (asymmetric bin ranges between the two
    directions: K2's hints must be validated), noise levels and densities; pruned + hinted search == every-SAD search."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tools"))
    import stress_match
    rng = np.random.default_rng(2024)
    assert all([stress_match.run_case(rng, i) for i in range(16)])


def test_bench_json_contract():
    """bench.py prints ONE JSON line with the keys the driver reads (tools/check_bench_contract.py; small batch, no CPU leg)."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tools"))
    import check_bench_contract
    check_bench_contract.main()
