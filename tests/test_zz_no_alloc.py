"""vo_bucket_matches_noalloc (no counterpart in the reference): the bucketing
kernel with its scratch in a static device array of the library, for callers that may not allocate.  Called through the C ABI
with device pointers and compared with the existing entry point vo_bucket_matches on the same lists -- integer / copied
records, so the tolerance is the suite's: bit-exact.  Two calls in a row on DIFFERENT lists (the array keeps the previous
call's contents), then the first list again.  Collected last on purpose (file name), and self-contained: it passes alone in a
fresh process and after the other tests."""
import numpy as np
import pytest

from hls_final_visual_odometry_b200 import Engine
from hls_final_visual_odometry_b200.api import P_MATCH_DTYPE, VoError

pytestmark = pytest.mark.gpu


def synthetic_matches(n, w, h, seed):
    """a raw match list as K3 emits it: integer coordinates, indices -1 (src/myMatch.cpp:189)"""
    rng = np.random.default_rng(seed)
    m = np.zeros(n, dtype=P_MATCH_DTYPE)
    names = P_MATCH_DTYPE.names
    for k in names:
        m[k] = -1
    u = rng.integers(9, w - 9, n)
    v = rng.integers(9, h - 9, n)
    m[names[6]] = u  # u1c
    m[names[7]] = v  # v1c
    m[names[0]] = np.clip(u + rng.integers(-40, 41, n), 0, w - 1)  # u1p
    m[names[1]] = np.clip(v + rng.integers(-20, 21, n), 0, h - 1)  # v1p
    return m


def run_noalloc(eng, torch, matches, mf, bw, bh, out_cap):
    raw = np.ascontiguousarray(matches).view(np.uint8).reshape(-1)
    d_in = torch.from_numpy(raw.copy()).cuda() if len(raw) else torch.zeros(48, dtype=torch.uint8, device="cuda")
    d_out = torch.full((max(out_cap, 1) * 48,), 0xAB, dtype=torch.uint8, device="cuda")
    d_res = torch.full((2,), 77, dtype=torch.int32, device="cuda")  # garbage: the call itself must reset it
    torch.cuda.synchronize()
    eng.bucket_matches_noalloc(d_in.data_ptr(), len(matches), mf, bw, bh, d_out.data_ptr(), out_cap, d_res.data_ptr())
    eng.synchronize()
    n, status = (int(x) for x in d_res.cpu().numpy())
    got = d_out.cpu().numpy()[: n * 48].view(P_MATCH_DTYPE).copy()
    return got, n, status


def test_noalloc_bucketing_equals_the_existing_entry_point_on_two_calls_in_a_row():
    torch = pytest.importorskip("torch")
    w, h = 1241, 376
    eng = Engine(w, h, max_frames=2, max_pairs=1, device=0)
    try:
        lists = [synthetic_matches(17000, w, h, 1), synthetic_matches(3100, w // 2, h // 2, 2), synthetic_matches(17000, w, h, 1),
                 synthetic_matches(6, w, h, 3), synthetic_matches(0, w, h, 4)]
        params = [(2, 50.0, 50.0), (3, 37.5, 61.0), (2, 50.0, 50.0), (2, 50.0, 50.0), (2, 50.0, 50.0)]
        for i, (m, (mf, bw, bh)) in enumerate(zip(lists, params)):
            want = eng.bucket_matches(m, mf, bw, bh)
            got, n, status = run_noalloc(eng, torch, m, mf, bw, bh, out_cap=max(len(m), 1))
            assert status == 0, f"call {i}: device status {status}"
            assert n == len(want), f"call {i}: {n} bucketed matches, existing entry point {len(want)}"
            assert got.tobytes() == want.tobytes(), f"call {i}: records differ"
        # capacity of the caller's output: flagged on the device, nothing written past it
        want = eng.bucket_matches(lists[0], 2, 50.0, 50.0)
        got, n, status = run_noalloc(eng, torch, lists[0], 2, 50.0, 50.0, out_cap=len(want) - 1)
        assert (n, status) == (0, 3)
        # and the next call is unaffected by the failed one
        got, n, status = run_noalloc(eng, torch, lists[1], 3, 37.5, 61.0, out_cap=len(lists[1]))
        assert status == 0 and got.tobytes() == eng.bucket_matches(lists[1], 3, 37.5, 61.0).tobytes()
    finally:
        eng.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    torch = pytest.importorskip("torch")
    eng = Engine(1241, 376, max_frames=2, max_pairs=1, device=0)
    try:
        buf = torch.zeros(48 * 8, dtype=torch.uint8, device="cuda")
        res = torch.zeros(2, dtype=torch.int32, device="cuda")
        with pytest.raises(VoError) as ei:
            eng.bucket_matches_noalloc(buf.data_ptr(), 32769, 2, 50.0, 50.0, buf.data_ptr(), 8, res.data_ptr())
        assert "limit of the static scratch" in str(ei.value)
        with pytest.raises(VoError):  # misaligned list
            eng.bucket_matches_noalloc(buf.data_ptr() + 4, 4, 2, 50.0, 50.0, buf.data_ptr(), 8, res.data_ptr())
    finally:
        eng.close()
