"""Diagnostic: our CUDA path vs the compiled reference on the same inputs.
Run on a GPU box: python tools/gpu_compare.py [P W H seed] ...  Prints per-stage
mismatch counts and simple CUDA-event timings.  (Test infrastructure.)"""
import sys, os, time, json
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
from trase_b200 import synth, _raster
from oracle import ref_loader


def ours_forward(sc, bg, **kw):
    cam = sc["cam"]
    e = torch.Tensor([])
    res = _raster.rasterize_forward(bg, sc["means3D"], e, sc["opacities"], sc["scales"], sc["rotations"], 1.0, e,
                                    cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"], cam["H"],
                                    cam["W"], sc["shs"], sc["sh_objs"], sc["sh_degree"], cam["campos"], False,
                                    kw.get("debug", False))
    return res


def ours_backward(sc, bg, res, g):
    cam = sc["cam"]
    e = torch.Tensor([])
    out = _raster.rasterize_backward(bg, sc["means3D"], res.radii, e, sc["scales"], sc["rotations"], 1.0, e,
                                     cam["viewmatrix"], cam["projmatrix"], cam["tanfovx"], cam["tanfovy"],
                                     g["dL_dcolor"], sc["shs"], g["dL_dobjects"], sc["sh_objs"], sc["sh_degree"],
                                     cam["campos"], res.geom, res.num_rendered, res.binning, res.img, False,
                                     capacity=res.capacity)
    names = ["means2D", "colors_precomp", "objects", "opacities", "means3D", "cov3D", "sh", "scales", "rotations"]
    return dict(zip(names, out))


def timeit(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(True), torch.cuda.Event(True)) for _ in range(n)]
    for a, b in ev:
        a.record(); fn(); b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    return ts[len(ts) // 2]


def compare(P, W, H, seed, bgval=0.0, timing=True):
    dev = torch.device("cuda:0")
    sc = synth.scene_to(synth.make_scene(P, W, H, seed), dev)
    g = {k: v.to(dev) for k, v in synth.make_upstream_grads(W, H, seed).items()}
    bg = torch.full((3,), bgval, device=dev)
    ref = ref_loader.run_forward(sc, bg)
    res = ours_forward(sc, bg)
    torch.cuda.synchronize()
    T = ((W + 15) // 16) * ((H + 15) // 16)
    v = _raster.workspace_views(res.geom, res.img, res.binning, P, W, H, res.capacity, res.num_rendered)
    rep = dict(P=P, W=W, H=H, seed=seed, bg=bgval, R_ref=int(ref["num_rendered"]), R=res.num_rendered)
    vis = ref["radii"] > 0
    rep["visible"] = int(vis.sum())
    rep["radii_mismatch"] = int((ref["radii"] != res.radii).sum())
    rep["tiles_mismatch"] = int((ref["tiles_touched"] != v["tiles_touched"]).sum())
    for name, a, b in [("depths", ref["depths"], v["depths"]), ("means2D", ref["means2D"], v["means2D"]),
                       ("conic_opacity", ref["conic_opacity"], v["conic_opacity"]),
                       ("rgb", ref["rgb"], v["rgbd"][:, :3]), ("cov3D", ref["cov3D"], v["cov3D"])]:
        m = vis if a.dim() == 1 else vis[:, None].expand_as(a)
        rep[name + "_neq"] = int(((a != b) & m).sum())
        rep[name + "_maxabs"] = float(((a - b).abs() * m).max()) if P else 0.0
    if rep["R_ref"] == rep["R"]:
        rep["keys_mismatch"] = int((ref["point_list_keys"] != v["point_list_keys"]).sum())
        rep["list_mismatch"] = int((ref["point_list"] != v["point_list"]).sum())
    rep["ranges_mismatch"] = int((ref["ranges"] != v["ranges"]).sum())
    rep["n_contrib_mismatch"] = int((ref["n_contrib"] != v["n_contrib"]).sum())
    rep["final_T_maxabs"] = float((ref["accum_alpha"] - v["final_T"]).abs().max())
    for name, a, b in [("color", ref["color"], res.color), ("objects", ref["objects"], res.objects),
                       ("depth", ref["depth"], res.depth)]:
        rep[name + "_maxabs"] = float((a - b).abs().max())
    # backward: reference mean of 3 runs
    rb = [ref_loader.run_backward(sc, bg, ref, g["dL_dcolor"], g["dL_dobjects"]) for _ in range(3)]
    ob = ours_backward(sc, bg, res, g)
    torch.cuda.synchronize()
    for k in ob:
        r = sum(x[k].double() for x in rb) / 3
        o = ob[k].double().view_as(r)
        denom = r.abs().max().clamp_min(1e-30)
        rep["grad_" + k + "_relmax"] = float((r - o).abs().max() / denom)
        rep["grad_" + k + "_relnorm"] = float((r - o).norm() / r.norm().clamp_min(1e-30))
    if timing:
        rep["t_ref_fwd_ms"] = timeit(lambda: ref_loader.run_forward(sc, bg))
        rep["t_our_fwd_ms"] = timeit(lambda: ours_forward(sc, bg))
        rep["t_ref_bwd_ms"] = timeit(lambda: ref_loader.run_backward(sc, bg, ref, g["dL_dcolor"], g["dL_dobjects"]))
        rep["t_our_bwd_ms"] = timeit(lambda: ours_backward(sc, bg, res, g))
    return rep


if __name__ == "__main__":
    cases = [(20000, 128, 128, 1, 0.0), (20000, 128, 128, 2, 1.0), (50000, 200, 136, 3, 0.0),
             (300000, 1352, 1014, 2, 0.0)]
    if len(sys.argv) > 1:
        a = [float(x) for x in sys.argv[1:]]
        cases = [(int(a[0]), int(a[1]), int(a[2]), int(a[3]), a[4] if len(a) > 4 else 0.0)]
    out = []
    for c in cases:
        rep = compare(*c)
        print(json.dumps(rep), flush=True)
        out.append(rep)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/compare.json", "w"), indent=1)
