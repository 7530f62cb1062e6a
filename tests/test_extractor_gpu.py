"""Parity of the CUDA extractor (through the C ABI) with the CPU oracle, stage by stage and end to end.

Bars (BASELINE.json north_star): pyramid, blurred levels, keypoint sets (x, y, octave, response) and descriptors
bit-exact; angles bit-exact (tolerance 0, stated here), all on the same seeded synthetic frames.
"""
import numpy as np
import pytest

from oracle import oracle_lib as O
from orb_slam_cuda_b200 import synth
import orb_slam_cuda_b200 as X

pytestmark = pytest.mark.gpu

CONFIGS = {  # reference YAMLs: Examples/Monocular/KITTI00-02.yaml, TUM1.yaml, Examples/Stereo/EuRoC.yaml
    "kitti": dict(nfeatures=2000, sf=1.2, nl=8, ini=20, mn=7),
    "tum": dict(nfeatures=1000, sf=1.2, nl=8, ini=20, mn=7),
    "euroc": dict(nfeatures=1200, sf=1.2, nl=8, ini=20, mn=7),
}


def make(kind):
    c = CONFIGS[kind]
    return (X.ORBextractor(c["nfeatures"], c["sf"], c["nl"], c["ini"], c["mn"]),
            O.Extractor(c["nfeatures"], c["sf"], c["nl"], c["ini"], c["mn"]))


def kp_tuple(k):
    return [(float(a["x"]), float(a["y"]), float(a["response"])) for a in k]


@pytest.fixture(scope="module", params=["kitti", "tum", "euroc"])
def run(request):
    kind = request.param
    gx, ox = make(kind)
    img = synth.frame(kind, 3)
    kps, desc = gx(img)
    okps, odesc = ox.extract(img)
    return dict(kind=kind, gx=gx, ox=ox, img=img, kps=kps, desc=desc, okps=okps, odesc=odesc)


def test_tables(run):
    t = run["ox"].tables()
    gx = run["gx"]
    assert np.array_equal(gx.GetScaleFactors(), t["scale"])
    assert np.array_equal(gx.GetInverseScaleFactors(), t["inv_scale"])
    assert np.array_equal(gx.GetScaleSigmaSquares(), t["sigma2"])
    assert np.array_equal(gx.GetInverseScaleSigmaSquares(), t["inv_sigma2"])
    assert np.array_equal(gx.features_per_level(), t["quota"])
    assert np.array_equal(gx.umax(), t["umax"])


def test_pyramid_bit_exact(run):
    gx, ox = run["gx"], run["ox"]
    for l in range(gx.nlevels):
        assert gx.level_size(l) == ox.level_size(l)
        a, b = gx.image_pyramid_level(l, padded=True), ox.level_padded(l)
        assert a.shape == b.shape
        assert np.array_equal(a, b), f"level {l}: {(a != b).sum()} of {a.size} padded pixels differ"


def test_blur_bit_exact(run):
    gx, ox = run["gx"], run["ox"]
    for l in range(gx.nlevels):
        b = ox.level_blurred(l)
        if b is None:
            continue
        a = gx.debug_blurred_level(l)
        assert np.array_equal(a, b), f"level {l}: {(a != b).sum()} blurred pixels differ"


def test_fast_candidates(run):
    gx, ox = run["gx"], run["ox"]
    for l in range(gx.nlevels):
        a, b = gx.debug_candidates(l), ox.level_candidates(l)
        assert kp_tuple(a) == kp_tuple(b), f"level {l}: {len(a)} vs {len(b)} candidates"


def test_level_keypoints_and_angles(run):
    gx, ox = run["gx"], run["ox"]
    for l in range(gx.nlevels):
        a, b = gx.debug_level_keypoints(l), ox.level_keypoints(l)
        assert kp_tuple(a) == kp_tuple(b), f"level {l}: quadtree output differs ({len(a)} vs {len(b)})"
        assert np.array_equal(a["angle"].view(np.uint32), b["angle"].view(np.uint32)), f"level {l}: angle bits differ"
        assert np.array_equal(a["size"], b["size"]) and np.array_equal(a["octave"], b["octave"])


def test_extract_bit_exact(run):
    k, d, ok, od = run["kps"], run["desc"], run["okps"], run["odesc"]
    assert len(k) == len(ok)
    for f in KPF:
        assert np.array_equal(k[f].view(np.uint32), ok[f].view(np.uint32)), f"field {f} differs"
    assert np.array_equal(d, od), f"{np.unpackbits(d ^ od).sum()} descriptor bits differ"


KPF = ["x", "y", "size", "angle", "response", "octave", "class_id"]


def test_batch_equals_single():
    gx, ox = make("tum")
    frames = synth.frame_batch("tum", 5, pool=5, first=10)
    kps, desc, counts = gx.extract_batch(frames)
    for f in range(frames.shape[0]):
        ok, od = ox.extract(frames[f])
        n = counts[f]
        assert n == len(ok)
        assert np.array_equal(kps[f, :n].view(np.uint8), ok.view(np.uint8))
        assert np.array_equal(desc[f, :n], od)


@pytest.mark.parametrize("kind", ["kitti", "euroc"])
def test_single_frame_kernel_shapes(kind):
    """Batches of 1 and 2 frames run the single-frame launch sequence (programmatic dependent launch, resize tiles shared by six
    CTAs, four warps per blur tile, one keypoint per describe warp, 512-thread quadtree CTAs); a batch of 3 runs the throughput
    shapes.  All must give the oracle's bytes, on the host-array path, on repeated calls (graph replay) and on device buffers."""
    import torch
    gx, ox = make(kind)
    frames = synth.frame_batch(kind, 3, pool=3, first=20)
    want = [ox.extract(frames[f]) for f in range(3)]
    for b in (1, 2, 3, 2, 1):
        kps, desc, counts = gx.extract_batch(frames[:b])
        for f in range(b):
            n = counts[f]
            assert n == len(want[f][0])
            assert np.array_equal(kps[f, :n].view(np.uint8), want[f][0].view(np.uint8)) and np.array_equal(desc[f, :n], want[f][1])
    for rep in range(4):  # second call captures the graph, later ones replay it
        for f in (0, 1):
            k, d = gx(frames[f])
            assert np.array_equal(k.view(np.uint8), want[f][0].view(np.uint8)) and np.array_equal(d, want[f][1])
    cap = gx.default_cap()
    h, w = frames.shape[1:]
    for b in (2, 1):
        d_frames = torch.from_numpy(frames[:b]).cuda()
        d_kps = torch.zeros((b, cap, 7), dtype=torch.int32, device="cuda")
        d_desc = torch.zeros((b, cap, 32), dtype=torch.uint8, device="cuda")
        d_counts = torch.zeros(b, dtype=torch.int32, device="cuda")
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            for rep in range(3):  # back to back on one stream: a call's first kernel waits for the previous call's last
                gx.extract_batch_device(d_frames.data_ptr(), b, w, h, d_kps.data_ptr(), d_desc.data_ptr(), cap, d_counts.data_ptr(), 0, st.cuda_stream)
        st.synchronize()
        for f in range(b):
            n = int(d_counts[f])
            assert n == len(want[f][0])
            assert np.array_equal(d_kps[f, :n].cpu().numpy().view(np.uint8).reshape(-1), want[f][0].view(np.uint8).reshape(-1))
            assert np.array_equal(d_desc[f, :n].cpu().numpy(), want[f][1])


@pytest.mark.parametrize("pinned", [False, True])
def test_chunked_host_batch(pinned):
    """Host batches of >= 256 frames are pipelined in chunks of about 128 over three streams (H2D | kernels | D2H): every frame
    must come out as if extracted alone, and the pyramid accessor serves the last chunk by batch index."""
    import torch
    h, w, nf, b = 120, 160, 300, 259
    gx, ox = X.ORBextractor(nf, 1.2, 8, 20, 7), O.Extractor(nf, 1.2, 8, 20, 7)
    pool = np.stack([synth.textured_frame(900 + i, h, w) for i in range(7)])
    frames = pool[np.arange(b) % 7].copy()
    frames[:, 0, 0] = (np.arange(b) % 256).astype(np.uint8)  # make every frame distinct
    frames[:, 0, 1] = (np.arange(b) // 256).astype(np.uint8)
    cap = gx.default_cap()
    if pinned:
        tf = torch.from_numpy(frames).pin_memory()
        tk = torch.empty((b, cap, 7), dtype=torch.int32).pin_memory()
        td = torch.empty((b, cap, 32), dtype=torch.uint8).pin_memory()
        tc = torch.empty(b, dtype=torch.int32).pin_memory()
        kps, desc, counts = gx.extract_batch(tf.numpy(), tk.numpy().view(X.KP_DTYPE).reshape(b, cap), td.numpy(), tc.numpy(), cap)
    else:
        kps, desc, counts = gx.extract_batch(frames)
    for f in [0, 1, 63, 64, 127, 128, 129, 130, 131, 257, 258]:
        ok, od = ox.extract(frames[f])
        n = counts[f]
        assert n == len(ok), f
        assert np.array_equal(kps[f, :n].view(np.uint8), ok.view(np.uint8)), f
        assert np.array_equal(desc[f, :n], od), f
    ox.extract(frames[b - 1])
    assert np.array_equal(gx.image_pyramid_level(3, frame=b - 1, padded=True), ox.level_padded(3))
    with pytest.raises(Exception):
        gx.image_pyramid_level(0, frame=0)  # not in the last chunk any more


def test_async_host_batches_two_in_flight():
    """orbx_extract_batch_async / orbx_extract_batch_wait: three different batches enqueued back to back, two in flight at
    any time (their copies and kernels overlap on the device); every batch must equal the blocking call's result and —
    for sample frames — the oracle; the misuse cases fail cleanly."""
    import torch
    h, w, nf, b = 120, 160, 300, 150
    gx, ox = X.ORBextractor(nf, 1.2, 8, 20, 7), O.Extractor(nf, 1.2, 8, 20, 7)
    cap = gx.default_cap()
    pool = np.stack([synth.textured_frame(940 + i, h, w) for i in range(9)])

    def batch(k):
        fr = pool[(np.arange(b) + k) % 9].copy()
        fr[:, 0, 0] = (np.arange(b) * 3 + k) % 251
        return torch.from_numpy(fr).pin_memory()

    def outputs():
        tk = torch.empty((b, cap, 7), dtype=torch.int32).pin_memory()
        td = torch.empty((b, cap, 32), dtype=torch.uint8).pin_memory()
        tc = torch.empty(b, dtype=torch.int32).pin_memory()
        return tk.numpy().view(X.KP_DTYPE).reshape(b, cap), td.numpy(), tc.numpy(), (tk, td, tc)

    ins = [batch(k) for k in range(3)]
    ref = [gx.extract_batch(t.numpy().copy()) for t in ins]  # blocking, pageable
    outs = [outputs() for _ in range(3)]
    gx.extract_batch_async(ins[0].numpy(), *outs[0][:3], cap)
    gx.extract_batch_async(ins[1].numpy(), *outs[1][:3], cap)
    with pytest.raises(X.OrbxError):  # a third call in flight is refused
        gx.extract_batch_async(ins[2].numpy(), *outs[2][:3], cap)
    with pytest.raises(X.OrbxError):  # and so is a blocking call while asynchronous ones are pending
        gx.extract_batch(ins[2].numpy())
    gx.extract_batch_wait()
    gx.extract_batch_async(ins[2].numpy(), *outs[2][:3], cap)
    gx.extract_batch_wait()
    gx.extract_batch_wait()
    with pytest.raises(X.OrbxError):
        gx.extract_batch_wait()  # nothing left
    with pytest.raises(X.OrbxError):  # pageable buffers cannot be used asynchronously
        gx.extract_batch_async(ins[0].numpy().copy(), *outs[0][:3], cap)
    for k in range(3):
        kps, desc, counts = outs[k][:3]
        rk, rd, rc = ref[k]
        assert np.array_equal(counts, rc), k
        for f in range(b):
            n = counts[f]
            assert np.array_equal(kps[f, :n].view(np.uint8), rk[f, :n].view(np.uint8)), (k, f)
            assert np.array_equal(desc[f, :n], rd[f, :n]), (k, f)
        for f in (0, 77, b - 1):
            ok, od = ox.extract(ins[k].numpy()[f])
            n = counts[f]
            assert n == len(ok) and np.array_equal(kps[f, :n].view(np.uint8), ok.view(np.uint8)) and np.array_equal(desc[f, :n], od), (k, f)
    kps, desc, counts = gx.extract_batch(ins[0].numpy())  # the blocking call works again afterwards
    assert np.array_equal(counts, ref[0][2])


@pytest.mark.parametrize("kind,nfeatures", [("kitti", 4000), ("tum", 2000)])
def test_monocular_initialisation_extractor(kind, nfeatures):
    """Tracking builds a second extractor with 2 * nFeatures for the monocular initialisation (src/Tracking.cc:133:
    mpIniORBextractor = new ORBextractor(2*nFeatures, ...)): 4000 features for KITTI, 2000 for TUM.  Bigger per-level quotas mean
    deeper quadtrees and more shared memory in the quadtree kernel; the result must still equal the oracle bit for bit, single and
    batched."""
    gx, ox = X.ORBextractor(nfeatures, 1.2, 8, 20, 7), O.Extractor(nfeatures, 1.2, 8, 20, 7)
    assert np.array_equal(gx.features_per_level(), ox.tables()["quota"])
    frames = np.stack([synth.frame(kind, 70 + i) for i in range(3)])
    kps, desc, counts = gx.extract_batch(frames)
    for f in range(3):
        ok, od = ox.extract(frames[f])
        n = counts[f]
        assert n == len(ok) and n > nfeatures * 0.8
        assert np.array_equal(kps[f, :n].view(np.uint8), ok.view(np.uint8))
        assert np.array_equal(desc[f, :n], od)
    k1, d1 = gx(frames[1])
    ok, od = ox.extract(frames[1])
    assert np.array_equal(k1.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d1, od)


def test_pattern_variants():
    """bit_pattern_31_ as the reference compiles it (default: element 96 = VX_FAILURE - 2 = -3, src/ORBextractor.cc:261-262) and the
    canonical ORB table: both equal the oracle's variant bit for bit, they differ from each other in bit 0 of descriptor byte 3 only
    (pair 24), keypoints are the same, and switching back and forth (incl. the replayed single-frame graph) takes effect."""
    gx, ox = make("tum")
    img = synth.frame("tum", 21)
    try:
        out = {}
        for variant in (X.ORBextractor.PATTERN_REFERENCE, X.ORBextractor.PATTERN_CANONICAL, X.ORBextractor.PATTERN_REFERENCE):
            gx.set_pattern(variant)
            O.set_pattern_variant(variant)
            for _ in range(3):  # the third call replays the captured graph
                k, d = gx(img)
            ok, od = ox.extract(img)
            assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od), variant
            out[variant] = (k.copy(), d.copy())
        (k0, d0), (k1, d1) = out[0], out[1]
        assert np.array_equal(k0.view(np.uint8), k1.view(np.uint8))
        diff = d0 ^ d1
        assert diff.any() and not np.delete(diff, 3, axis=1).any() and not (diff[:, 3] & 0xFE).any()
    finally:
        O.set_pattern_variant(O.PATTERN_REFERENCE)


def test_strided_input_and_empty():
    gx, ox = make("tum")
    big = np.zeros((500, 700), np.uint8)
    img = synth.frame("tum", 20)
    big[5:485, 30:670] = img
    k, d = gx(big[5:485, 30:670])
    ok, od = ox.extract(img)
    assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od)
    k, d = gx(np.empty((0, 0), np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)


def test_small_and_odd_sizes():
    for (h, w, nf) in [(120, 160, 300), (97, 211, 200), (240, 320, 500)]:
        gx, ox = X.ORBextractor(nf, 1.2, 8, 20, 7), O.Extractor(nf, 1.2, 8, 20, 7)
        img = synth.textured_frame(4242 + h, h, w)
        k, d = gx(img)
        ok, od = ox.extract(img)
        assert len(k) == len(ok), (h, w)
        assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od), (h, w)


def test_repeated_single_frames_graph_replay():
    """The second single-frame call of a geometry captures a CUDA graph, later calls replay it: every frame of a stream
    must still equal the oracle, mvImagePyramid must follow the current frame, and a size change must re-capture."""
    gx, ox = make("tum")
    for i in range(6):
        img = synth.frame("tum", 40 + i)
        k, d = gx(img)
        ok, od = ox.extract(img)
        assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od), i
        assert np.array_equal(gx.image_pyramid_level(2, padded=True), ox.level_padded(2)), i
    for i in range(4):  # other geometry on the same handle, then back
        img = synth.textured_frame(700 + i, 240, 320)
        k, d = gx(img)
        ok, od = ox.extract(img)
        assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od), i
    img = synth.frame("tum", 77)
    k, d = gx(img)
    ok, od = ox.extract(img)
    assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od)


@pytest.mark.parametrize("amp", [255, 120])
def test_dense_noise_image(amp):
    """White noise: most pixels pass the FAST quick test, so the optimistic single-chunk survivor list of a strip
    overflows and the pass is redone in row chunks; cells hold far more corners than a natural image gives."""
    rng = np.random.default_rng(5)
    img = (128 + (rng.random((240, 320)) - 0.5) * amp).astype(np.uint8)
    gx, ox = X.ORBextractor(500, 1.2, 8, 20, 7), O.Extractor(500, 1.2, 8, 20, 7)
    k, d = gx(img)
    ok, od = ox.extract(img)
    for l in range(8):
        a, b = gx.debug_candidates(l), ox.level_candidates(l)
        assert kp_tuple(a) == kp_tuple(b), f"level {l}: {len(a)} vs {len(b)} candidates"
    assert len(k) == len(ok)
    assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od)


def test_full_range_noise_cannot_overflow_a_cell():
    """uniform 0..255 noise: up to ~200 corners survive NMS in a 31 x 32 cell (the r02 build had 128 slots per cell and
    returned ORBX_E_OVERFLOW here); slots are now sized by the NMS packing bound ceil(w/2) * ceil(h/2) of the cell"""
    rng = np.random.default_rng(15)
    for shape, nf in (((300, 420), 1500), ((376, 1241), 2000), ((120, 150), 300)):
        img = rng.integers(0, 256, shape, dtype=np.uint8)
        gx, ox = X.ORBextractor(nf, 1.2, 8, 20, 7), O.Extractor(nf, 1.2, 8, 20, 7)
        k, d = gx(img)
        ok, od = ox.extract(img)
        for l in range(8):
            a, b = gx.debug_candidates(l), ox.level_candidates(l)
            assert kp_tuple(a) == kp_tuple(b), f"level {l}: {len(a)} vs {len(b)} candidates"
        assert len(k) == len(ok)
        assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od)
    # bright pixels on every second row and column: a periodic pattern whose corners appear in the resized levels only
    img = np.zeros((200, 260), np.uint8)
    img[::2, ::2] = 255
    gx, ox = X.ORBextractor(1000, 1.2, 8, 20, 7), O.Extractor(1000, 1.2, 8, 20, 7)
    k, d = gx(img)
    ok, od = ox.extract(img)
    assert len(k) == len(ok) and np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od)


def test_quadtree_standalone_random():
    rng = np.random.default_rng(7)
    for trial in range(40):
        W, H = int(rng.integers(60, 1300)), int(rng.integers(60, 480))
        n, N = int(rng.integers(1, 3000)), int(rng.integers(1, 500))
        pts = np.unique(np.stack([rng.integers(0, H, n), rng.integers(0, W, n)], 1), axis=0)
        if trial % 2:
            rng.shuffle(pts)
        keys = np.zeros(len(pts), X.KP_DTYPE)
        keys["x"], keys["y"] = pts[:, 1], pts[:, 0]
        keys["response"] = rng.integers(7, 60, len(pts))
        a = X.distribute_octtree(keys, 16, 16 + W, 16, 16 + H, N)
        b = O.distribute(keys, 16, 16 + W, 16, 16 + H, N)
        assert kp_tuple(a) == kp_tuple(b), f"trial {trial}: W={W} H={H} n={len(pts)} N={N}: {len(a)} vs {len(b)}"


@pytest.mark.parametrize("sf,nl", [(2.0, 4), (1.5, 5), (2.5, 3), (1.1, 8)])
def test_other_scale_factors(sf, nl):
    # 2.0 = cv::resize's INTER_AREA shortcut on even sizes; 2.5 = source windows wider than 8 bytes (byte-gather kernel)
    gx, ox = X.ORBextractor(800, sf, nl, 20, 7), O.Extractor(800, sf, nl, 20, 7)
    img = synth.frame("tum", 61)
    k, d = gx(img)
    ok, od = ox.extract(img)
    for l in range(nl):
        a, b = gx.image_pyramid_level(l, padded=True), ox.level_padded(l)
        assert np.array_equal(a, b), f"sf {sf} level {l}: {(a != b).sum()} padded pixels differ"
    assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od)


def test_full_size_batch_properties():
    """BASELINE.json's full size (512 KITTI frames of 1241x376, 2000 features per frame) through the chunked host pipeline, checked by
    properties that do not need the oracle on every frame: copies of the same frame anywhere in the batch give byte-identical
    keypoints and descriptors, a second run of the batch reproduces the first (no state leaks between chunks, lanes or runs), every
    frame holds nfeatures..nfeatures+3 keypoints in level-major order inside the image, and the distinct frames equal the oracle."""
    nf, b, pool = 2000, 512, 8
    gx, ox = X.ORBextractor(nf, 1.2, 8, 20, 7), O.Extractor(nf, 1.2, 8, 20, 7)
    base = np.stack([synth.frame("kitti", 300 + i) for i in range(pool)])
    frames = base[np.arange(b) % pool].copy()
    kps, desc, counts = gx.extract_batch(frames)
    kps, desc, counts = kps.copy(), desc.copy(), counts.copy()
    kps2, desc2, counts2 = gx.extract_batch(frames)
    assert np.array_equal(counts, counts2)
    for f in range(b):  # slots beyond counts[f] are not written
        n = counts[f]
        assert np.array_equal(kps[f, :n].view(np.uint8), kps2[f, :n].view(np.uint8)) and np.array_equal(desc[f, :n], desc2[f, :n]), f
    first_of = {}
    for f in range(b):
        key = f % pool
        n = counts[f]
        assert nf <= n <= nf + 8 * 3, (f, n)   # the quadtree may overshoot by up to 3 per level
        k = kps[f, :n]
        assert (np.diff(k["octave"]) >= 0).all() and k["x"].min() >= 0 and k["x"].max() < 1241 and k["y"].min() >= 0 and k["y"].max() < 376
        if key in first_of:
            g = first_of[key]
            assert counts[g] == n and np.array_equal(kps[g, :n].view(np.uint8), k.view(np.uint8)) and np.array_equal(desc[g, :n], desc[f, :n])
        else:
            first_of[key] = f
            ok, od = ox.extract(frames[f])
            assert n == len(ok) and np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(desc[f, :n], od)
    assert len(first_of) == pool


@pytest.mark.parametrize("name", ["kitti_0", "tum_1", "euroc_2", "levels4_sf15", "noise", "blocks", "dim", "few"])
def test_against_reference_object_code_golden(name):
    """The CUDA extractor against the REFERENCE's own compiled ORBextractor.cc (tests/golden/ref_extractor.npz, written by
    tools/gen_extractor_golden.py from oracle/_ref/libref_extractor.so): order, x, y, size, angle, response, octave and the
    32 descriptor bytes, tolerance 0 — no oracle in between."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(__file__))
    import ref_extractor_cases as RC
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_extractor.npz"))
    make_img, kw = RC.CASES[name]
    img = make_img()
    assert int(g[name + "/img_sum"]) == int(img.astype(np.int64).sum())
    a = RC.oracle_kwargs(kw)
    gx = X.ORBextractor(a["nfeatures"], a["scale_factor"], a["nlevels"], a["ini_th"], a["min_th"])
    k, d = gx(np.ascontiguousarray(img))
    rk = g[name + "/kps"]
    assert len(k) == len(rk) > 0
    for i, f in enumerate(("x", "y", "size", "angle", "response")):
        assert np.array_equal(k[f], rk[:, i]), f
    assert np.array_equal(k["octave"], rk[:, 5].astype(np.int32))
    assert np.array_equal(d, g[name + "/desc"])
