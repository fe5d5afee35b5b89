"""CPU suite: pins the C restatement (oracle/gsdf_oracle.c) against the golden vectors produced by the reference's
own OMP build, and — where oracle/_ref exists — against that build directly, bit for bit."""
import numpy as np
import pytest

from conftest import ROOT, load_golden

GOLDENS = ["grad_160x120_2cm_t5", "base_160x120_2cm_t5", "grad_160x120_1cm_t10"]


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def make_oracle(po, g):
    return po.Oracle(int(g["variant"]), g["voxel_size"], g["trunc"], int(g["width"]), int(g["height"]), g["K"])


@pytest.mark.parametrize("name", GOLDENS)
def test_restatement_matches_reference_golden(po, name):
    g = load_golden(name)
    o = make_oracle(po, g)
    assert np.array_equal(bits(o.normal_cache()), bits(g["normal_cache"]))
    n0 = o.normals(g["depth"][0])
    same = (bits(n0) == bits(g["normals0"])) | (np.isnan(n0) & np.isnan(g["normals0"]))
    assert same.all()
    n_inject = g["inject_R"].shape[0]
    for i in range(n_inject):
        R, t, upd, pts = o.update(g["depth"][i], g["poses"][i])
        assert np.array_equal(bits(R), bits(g["inject_R"][i])) and np.array_equal(bits(t), bits(g["inject_t"][i]))
        assert upd > 0 and pts > 0
        if i == 0:
            k, d, w, gr = o.download()
            assert np.array_equal(k, g["f0_keys"])
            assert np.array_equal(bits(d), bits(g["f0_dist"])) and np.array_equal(bits(w), bits(g["f0_weight"]))
            assert np.array_equal(bits(gr), bits(g["f0_grad"]))
    k, d, w, gr = o.download()
    assert np.array_equal(k, g["keys"])  # voxel key set bit-exact
    assert np.array_equal(bits(d), bits(g["dist"])) and np.array_equal(bits(w), bits(g["weight"]))
    assert np.array_equal(bits(gr), bits(g["grad"]))
    phi, grad, wq = o.query(g["q_pts"])
    assert np.array_equal(bits(phi), bits(g["q_phi"])) and np.array_equal(bits(grad), bits(g["q_grad"]))
    assert np.array_equal(bits(wq), bits(g["q_w"]))
    # one Gauss-Newton iteration: pose after exp(-xi) * pose within 1e-6 (LLT / sin / cos differ in the last ulp)
    o.set_opt(1, 1e-3, 1.0)
    o.set_pose(g["one_iter_start"])
    conv, _ = o.optimize(g["depth"][n_inject])
    assert conv == bool(g["one_iter_conv"])
    assert np.abs(o.get_pose() - g["one_iter_pose"]).max() < 2e-6


@pytest.mark.parametrize("name", GOLDENS[:2])
def test_restatement_free_running_matches_reference(po, name):
    """main_scan_3d.cpp:208-266 ordering: frame 0 fused at identity, then track; fuse only if converged."""
    g = load_golden(name)
    o = po.Oracle(int(g["variant"]), g["free_voxel_size"], g["trunc"], int(g["free_width"]), int(g["free_height"]),
                  g["free_K"])
    ident = np.array([0, 0, 0, 0, 0, 0, 1], np.float32)
    o.update(g["free_depth"][0], ident)
    o.set_pose(ident)
    for i in range(1, g["free_depth"].shape[0]):
        conv, it = o.optimize(g["free_depth"][i])
        assert conv == bool(g["free_conv"][i])
        if conv:
            assert it == int(g["free_iters"][i])
            o.update(g["free_depth"][i], o.get_pose())
        assert np.abs(o.get_pose() - g["free_traj"][i]).max() < 2e-4  # 0.2 mm: ulp-level solve differences move a few lookups
    k, d, w, gr = o.download()
    kg = g["free_keys"]
    # poses agree to ~1e-6 only, so single boundary voxels may differ: Jaccard, not equality
    a = {tuple(x) for x in k}
    b = {tuple(x) for x in kg}
    assert len(a & b) / len(a | b) > 0.999


def test_restatement_vs_reference_build_live(po, synth):
    """Where the reference build is present (this container; it also travels to the GPU box as a .so)."""
    if not po.have_ref():
        pytest.skip("oracle/_ref not built")
    W, H = 160, 120
    frames, poses, K, _ = synth.make_sequence(60, W, H, scene_seed=99, noise_seed=5)
    for variant, vs, tr in [(0, 0.016, 5.0), (1, 0.008, 5.0), (0, 0.004, 10.0)]:
        o = po.Oracle(variant, vs, tr, W, H, K)
        r = po.RefScan(variant, vs, tr, W, H, K)
        for i in range(3):
            o.update(frames[i], poses[i])
            r.update(frames[i], poses[i])
        ko, do, wo, go = o.download()
        kr, dr, wr, gr = r.download()
        assert np.array_equal(ko, kr)
        assert np.array_equal(bits(do), bits(dr)) and np.array_equal(bits(wo), bits(wr)) and np.array_equal(bits(go), bits(gr))
        rng = np.random.default_rng(3)
        idx = rng.integers(0, len(ko), 3000)
        pts = (ko[idx].astype(np.float32) + rng.uniform(-0.7, 0.7, (3000, 3)).astype(np.float32)) * np.float32(vs)
        for a, b in zip(o.query(pts), r.query(pts)):
            assert np.array_equal(bits(a), bits(b))
        o.set_pose(poses[2]); r.set_pose(poses[2])
        co, cr = o.optimize(frames[3]), r.optimize(frames[3])
        assert co[0] == cr[0]
        assert np.abs(o.get_pose() - r.get_pose()).max() < 2e-5


def test_pose_algebra_and_solve(po, pkg):
    """Host pose algebra of the product library (no GPU needed) against the oracle / reference (Sophus, Eigen LLT)."""
    rng = np.random.default_rng(0)
    o = po.Oracle(0, 0.02, 5.0, 32, 24, np.array([[50, 0, 16], [0, 50, 12], [0, 0, 1]], np.float32))
    refs = [o]
    if po.have_ref():
        refs.append(po.RefScan(0, 0.02, 5.0, 32, 24, o.K.reshape(3, 3)))
    for _ in range(200):
        q = rng.standard_normal(4).astype(np.float32)
        pose = np.concatenate([rng.uniform(-2, 2, 3), q]).astype(np.float32)
        R, pn = pkg.pose_matrix(pose)
        xi = (rng.standard_normal(6) * rng.choice([1e-7, 1e-3, 0.1])).astype(np.float32)
        out = pkg.pose_apply_increment(pn, xi)
        A = rng.standard_normal((12, 6)).astype(np.float32)
        Hm = (A.T @ A).astype(np.float32)
        gv = rng.standard_normal(6).astype(np.float32)
        x = pkg.llt_solve6(Hm, gv)
        for r in refs:
            Rr, pr = r.pose_matrix(pose)
            assert np.array_equal(bits(R), bits(Rr)) and np.array_equal(bits(pn), bits(pr))
            assert np.abs(out - r.apply_increment(pn, xi)).max() < 2e-6
            xr, _ = r.llt_solve(Hm, gv)
            assert np.abs(x - xr).max() <= 1e-4 * max(1.0, np.abs(xr).max())


def test_synth_generator_is_seeded(synth):
    a = synth.make_sequence(50, 64, 48)[0][:2]
    b = synth.make_sequence(50, 64, 48)[0][:2]
    assert np.array_equal(a, b)
    assert (a > 0).mean() > 0.05


def test_marching_cubes_case_table_equals_the_reference_builds(po):
    """The nibble-packed case table in csrc/gsdf_mesh.cuh, row by row, against the triTable inside the reference build
    (mesh/LayeredMarchingCubesNoColor.cpp:96-352), and the rule that replaces its edgeTable (:67-94)."""
    import ctypes as C
    import re
    if not po.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    text = (ROOT / "gradient-sdf_b200" / "csrc" / "gsdf_mesh.cuh").read_text()
    body = text[text.index("kTriTable[256] = {"):]
    body = body[:body.index("};")]
    rows = [int(x, 16) for x in re.findall(r"0x([0-9a-f]{16})ull", body)]
    assert len(rows) == 256
    L = C.CDLL(str(po.REF_SO))
    tri = (C.c_int * (256 * 16)).in_dll(L, "_ZN27LayeredMarchingCubesNoColor8triTableE")
    edge = (C.c_int * 256).in_dll(L, "_ZN27LayeredMarchingCubesNoColor9edgeTableE")
    corners = [(0, 1), (1, 2), (2, 3), (3, 0), (4, 5), (5, 6), (6, 7), (7, 4), (0, 4), (1, 5), (2, 6), (3, 7)]
    for c in range(256):
        ref = [tri[16 * c + i] for i in range(16)]
        n = ref.index(-1)
        assert rows[c] >> 60 == n and n % 3 == 0
        assert [(rows[c] >> (4 * i)) & 0xF for i in range(n)] == ref[:n]
        mask = sum(1 << e for e, (a, b) in enumerate(corners) if ((c >> a) & 1) != ((c >> b) & 1))
        assert mask == edge[c]
        assert all(mask >> e & 1 for e in ref[:n])  # a case's triangles only name edges that cross the surface
