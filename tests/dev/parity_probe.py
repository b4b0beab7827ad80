"""Development probe (runs on a GPU box): where do our trajectory / gradients part from the reference's own CUDA build?

    python tests/dev/parity_probe.py flips  <outdir>     ReLU-mask flips between the reference's fp16-accumulated hidden activations and
                                                        fp32 accumulation, and the grid-gradient agreement on the flip-free samples
    python tests/dev/parity_probe.py state  <outdir>     trainer state (master, Adam m / v / steps, gradients) of `ref_driver state` against
                                                        ours along the trajectory of curve_cfg1_2d_T19, per grid level
    python tests/dev/parity_probe.py curve  <outdir> <tag>   1000-step loss of OUR trainer (env switches / TCNN_B200_LIB select the variant)

Writes JSON summaries into <outdir>; the large dumps stay in /tmp on the box."""
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "tiny-cuda-nn_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)
from oracle import tcnn_oracle as O  # noqa: E402
from tests import common  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
FLIP_CASES = {
    "cfg1_2d_T19": (common.CONFIG_1, 2, 3, 1024),
    "shipped_2d_T15": (common.CONFIG_SHIPPED, 2, 3, 1024),
    "nerf_3d_T19": (common.CONFIG_NERF, 3, 4, 1024),
    "f4_3d_T14": (common.CONFIG_F4, 3, 3, 512),
}


def ref_dump(cfg, n_in, n_out, x, t, d, hidden=False):
    B = x.shape[0]
    x.tofile(os.path.join(d, "x.f32")); t.tofile(os.path.join(d, "t.f32"))
    with open(os.path.join(d, "cfg.json"), "w") as f:
        json.dump(cfg, f)
    env = dict(os.environ, REF_DUMP_HIDDEN="1") if hidden else None
    r = subprocess.run([REF, "dump", os.path.join(d, "cfg.json"), str(n_in), str(n_out), str(B), os.path.join(d, "x.f32"), os.path.join(d, "t.f32"), d],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    return json.load(open(os.path.join(d, "meta.json")))


def ours_grads(cfg, n_in, n_out, x_np, t_np):
    import torch
    import tinycudann as tcnn
    dev = torch.device("cuda:0")
    tr = tcnn.create_from_config(n_in, n_out, cfg, seed=1337, device=dev).trainer
    x, t = torch.from_numpy(x_np).to(dev), torch.from_numpy(t_np).to(dev)
    out = torch.empty(x.shape[0], 16, dtype=torch.half, device=dev)
    tr.training_step(x, t, run_optimizer=False, output=out)
    return tr.param_gradients.cpu().numpy(), tr.loss(), out.cpu().numpy()


def flips(outdir):
    """(a) how many ReLU masks differ between the reference (fp16 accumulation) and fp32 accumulation on the golden batch;
    (b) a batch made of flip-free samples only: grid-gradient rel-L2 of ours against the reference on it."""
    res = {}
    only = [n for n in os.environ.get("PROBE_CASES", "").split(",") if n]
    for name, (cfg, n_in, n_out, B) in FLIP_CASES.items():
        if only and name not in only:
            continue
        x, t = common.synth_inputs(4 * B, n_in, n_out, seed=1)   # 4 B candidates -> enough flip-free ones for a batch of B
        with tempfile.TemporaryDirectory() as d:
            meta = ref_dump(cfg, n_in, n_out, x, t, d, hidden=True)
            P, P_grid = meta["n_params"], meta["n_params_encoding"]
            P_mlp = P - P_grid
            W, H = cfg["network"]["n_neurons"], cfg["network"]["n_hidden_layers"]
            hid_ref = [np.fromfile(os.path.join(d, f"hidden_{k}.f16"), np.float16).reshape(4 * B, W) for k in range(H)]
            p16 = np.fromfile(os.path.join(d, "params_used.f16"), np.float16)
            level_scale = np.fromfile(os.path.join(d, "level_scale.f32"), np.float32)
            grads_all_ref = np.fromfile(os.path.join(d, "grads.f16"), np.float16)
            out_ref = np.fromfile(os.path.join(d, "output_padded.f16"), np.float16).reshape(4 * B, 16)
        g, m = common.oracle_configs(cfg, n_in, n_out, scales=level_scale)
        enc = common.oracle_encode(cfg, g, m, x, p16[m.n_params:])
        y, hid = O.mlp_forward(m, p16[: m.n_params], enc)
        flip_rows = np.zeros(4 * B, bool)
        n_flips, bias = [], []
        for k in range(H):
            a, b = hid_ref[k].astype(np.float32), np.asarray(hid[k], np.float32)
            fl = (a > 0) != (b > 0)
            n_flips.append(int(fl.sum()))
            flip_rows |= fl.any(axis=1)
            both = (a > 0) & (b > 0)
            bias.append(float((a[both] - b[both]).sum() / np.abs(b[both]).sum()))  # signed: < 0 means the reference's activations are smaller
        free = np.flatnonzero(~flip_rows)[:B]
        entry = {"B_probe": 4 * B, "flips_per_layer": n_flips, "flip_fraction": float(sum(n_flips)) / (4 * B * W * H),
                 "rows_with_flip": int(flip_rows.sum()), "hidden_signed_bias_ref_minus_fp32": bias, "flip_free_available": int((~flip_rows).sum())}
        # all 4B samples: ours against the reference
        ga, _, out_ours = ours_grads(cfg, n_in, n_out, x, t)
        entry["output_rel_l2_vs_ref"] = common.rel_l2(out_ours[:, :n_out], out_ref[:, :n_out])
        entry["output_bits_differ_vs_ref"] = int(np.count_nonzero(out_ours[:, :n_out].view(np.uint16) != out_ref[:, :n_out].view(np.uint16)))
        entry["output_rel_l2_oracle_fp32_vs_ref"] = common.rel_l2(np.asarray(y)[:, :n_out], out_ref[:, :n_out])
        nz = np.flatnonzero(grads_all_ref[P_mlp:].view(np.uint16) & 0x7FFF)
        entry["grid_rel_l2_all_samples"] = common.rel_l2(ga[P_mlp:][nz], grads_all_ref[P_mlp:][nz])
        entry["mlp_rel_l2_all_samples"] = common.rel_l2(ga[:P_mlp], grads_all_ref[:P_mlp])
        if len(free) == B:
            xs, ts = np.ascontiguousarray(x[free]), np.ascontiguousarray(t[free])
            with tempfile.TemporaryDirectory() as d:
                ref_dump(cfg, n_in, n_out, xs, ts, d)
                gr = np.fromfile(os.path.join(d, "grads.f16"), np.float16)
            go, _, _ = ours_grads(cfg, n_in, n_out, xs, ts)
            nz = np.flatnonzero(gr[P_mlp:].view(np.uint16) & 0x7FFF)
            entry["grid_rel_l2_flip_free"] = common.rel_l2(go[P_mlp:][nz], gr[P_mlp:][nz])
            entry["mlp_rel_l2_flip_free"] = common.rel_l2(go[:P_mlp], gr[:P_mlp])
            # oracle on the same batch (fp32 accumulation like ours)
            enc_s = common.oracle_encode(cfg, g, m, xs, p16[m.n_params:])
            ys, hs = O.mlp_forward(m, p16[: m.n_params], enc_s)
            lv, dy = O.loss_l2(ys, ts, n_out, relative=cfg["loss"]["otype"].lower() == "relativel2")
            dW, dx = O.mlp_backward(m, p16[: m.n_params], enc_s, hs, dy)
            gg = O.grid_backward(g, xs, dx[:, : g.n_features])
            entry["grid_rel_l2_flip_free_oracle_vs_ref"] = common.rel_l2(gg[nz], gr[P_mlp:][nz].astype(np.float64))
            np.savez_compressed(os.path.join(outdir, f"flipfree_{name}.npz"), config=json.dumps(cfg), n_in=n_in, n_out=n_out, B=B,
                                x=xs, t=ts, grads_mlp=gr[:P_mlp], grads_grid_nz_idx=nz.astype(np.uint32), grads_grid_nz_val=gr[P_mlp:][nz],
                                probe_B=4 * B, flips_per_layer=np.array(n_flips), rows_with_flip=int(flip_rows.sum()))
        res[name] = entry
        print(name, json.dumps(entry), flush=True)
    json.dump(res, open(os.path.join(outdir, "flips.json"), "w"), indent=1)


def level_slices(info):
    F, n_mlp = info.n_features_per_level, int(info.n_params_mlp)
    sl = {"mlp": slice(0, n_mlp)}
    for l in range(info.n_levels):
        sl[f"L{l}"] = slice(n_mlp + F * info.level_offset[l], n_mlp + F * info.level_offset[l + 1])
    return sl


def batches(B, n_in, n_out, steps, seed, cache):
    """The reference's pcg32 batches of `ref_driver curve/state`, cached as .npy (the numpy generator is the slow part)."""
    fx, ft = os.path.join(cache, f"x_{B}_{n_in}_{steps}_{seed}.npy"), os.path.join(cache, f"t_{B}_{n_in}_{n_out}_{steps}_{seed}.npy")
    if os.path.exists(fx) and os.path.exists(ft):
        return np.load(fx, mmap_mode="r"), np.load(ft, mmap_mode="r")
    rng = O.Pcg32(seed)
    X = np.empty((steps, B, n_in), np.float32); T = np.empty((steps, B, n_out), np.float32)
    for i in range(steps):
        X[i] = O.random_uniform(rng, B * n_in).reshape(B, n_in)
        T[i] = O.synth_target(X[i], n_out)
    np.save(fx, X); np.save(ft, T)
    return X, T


def state(outdir):
    import torch
    import tinycudann as tcnn
    g = np.load(os.path.join(ROOT, "tests", "golden", "curve_cfg1_2d_T19.npz"))
    cfg = json.loads(str(g["config"]))
    B, n_in, n_out, steps, seed = (int(g[k]) for k in ("B", "n_in", "n_out", "steps", "seed"))
    steps = int(os.environ.get("PROBE_STEPS", steps))
    dump_steps = [s for s in (1, 2, 3, 10, 30, 100, 300, 1000) if s <= steps]
    d = "/tmp/refstate"
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "cfg.json"), "w") as f:
        json.dump(cfg, f)
    t0 = time.time()
    r = subprocess.run([REF, "state", os.path.join(d, "cfg.json"), str(n_in), str(n_out), str(B), str(seed), d, str(steps)] + [str(s) for s in dump_steps],
                       capture_output=True, text=True, timeout=1800)
    print("ref_driver state:", r.returncode, r.stdout.strip(), r.stderr.strip()[-300:], f"{time.time() - t0:.0f}s", flush=True)
    assert r.returncode == 0
    ref_losses = np.fromfile(os.path.join(d, "losses.f32"), np.float32)
    X, T = batches(B, n_in, n_out, steps, seed, "/tmp")
    dev = torch.device("cuda:0")
    tr = tcnn.create_from_config(n_in, n_out, cfg, seed=1337, device=dev).trainer
    if os.environ.get("PROBE_UNFUSED"):
        tr.use_fused_step = False
    sl = level_slices(tr.info)
    init = tr.params_full_precision.cpu().numpy().copy()
    losses = torch.zeros(steps, dtype=torch.float32, device=dev)
    report = {"dump_steps": dump_steps, "levels": {}}
    # compact golden of the reference's state (tests/golden/state_cfg1_2d_T19.npz): every MLP weight + every 997th grid parameter
    n_mlp_, P_ = int(tr.info.n_params_mlp), int(tr.info.n_params)
    gidx = np.concatenate([np.arange(n_mlp_), np.arange(n_mlp_, P_, 997)]).astype(np.int64)
    golden = {"config": json.dumps(cfg), "B": B, "n_in": n_in, "n_out": n_out, "seed": seed, "index": gidx.astype(np.uint32), "n_params_mlp": n_mlp_,
              "golden_steps": np.array([s for s in dump_steps if s <= 100])}
    for i in range(steps):
        x, t = torch.from_numpy(np.ascontiguousarray(X[i])).to(dev), torch.from_numpy(np.ascontiguousarray(T[i])).to(dev)
        need = (i + 1) in dump_steps
        if need:  # separate calls so that the gradient buffer survives (the fused step re-zeroes it)
            tr.training_step(x, t, run_optimizer=False)
            grads = tr.param_gradients.cpu().numpy().copy()
            losses[i] = tr.loss_value[0]
            tr.optimizer_step()
        else:
            tr.training_step(x, t)
            losses[i] = tr.loss_value[0]
        if not need:
            continue
        s = i + 1
        rd = lambda name, dt: np.fromfile(os.path.join(d, f"s{s}_{name}"), dt)  # noqa: E731
        ref = {"master": rd("master.f32", np.float32), "grads": rd("grads.f16", np.float16), "m": rd("m.f32", np.float32), "v": rd("v.f32", np.float32),
               "steps": rd("steps.u32", np.uint32)}
        ours = {"master": tr.params_full_precision.cpu().numpy(), "grads": grads, "m": tr.first_moments.cpu().numpy(), "v": tr.second_moments.cpu().numpy(),
                "steps": tr.param_steps.cpu().numpy().view(np.uint32)}
        if s <= 100:
            for k_ in ("master", "m", "v", "steps", "grads"):
                golden[f"s{s}_{k_}"] = ref[k_][gidx]
            for name, rng_ in sl.items():  # per tensor: how many gradient entries are non-zero, mean step count (pins the zero-gradient skip)
                golden[f"s{s}_nz_{name}"] = int(np.count_nonzero(ref["grads"][rng_])); golden[f"s{s}_stepsum_{name}"] = int(ref["steps"][rng_].astype(np.int64).sum())
        rep = {}
        for name, rng_ in sl.items():
            go, gr = ours["grads"][rng_].astype(np.float64), ref["grads"][rng_].astype(np.float64)
            e = {
                "grad_rel_l2": common.rel_l2(go, gr),
                "grad_abs_ratio": float(np.abs(go).sum() / max(np.abs(gr).sum(), 1e-300)),       # < 1: our gradients are smaller in magnitude
                "grad_nz_ours": int(np.count_nonzero(go)), "grad_nz_ref": int(np.count_nonzero(gr)),
                "grad_nz_only_ours": int(np.count_nonzero((go != 0) & (gr == 0))), "grad_nz_only_ref": int(np.count_nonzero((go == 0) & (gr != 0))),
                "grad_cos": float((go * gr).sum() / max(np.linalg.norm(go) * np.linalg.norm(gr), 1e-300)),
                "delta_rel_l2": common.rel_l2(ours["master"][rng_] - init[rng_], ref["master"][rng_] - init[rng_]),
                "delta_norm_ratio": float(np.linalg.norm(ours["master"][rng_] - init[rng_]) / max(np.linalg.norm(ref["master"][rng_] - init[rng_]), 1e-300)),
                "m_rel_l2": common.rel_l2(ours["m"][rng_], ref["m"][rng_]), "v_rel_l2": common.rel_l2(ours["v"][rng_], ref["v"][rng_]),
                "v_sum_ratio": float(ours["v"][rng_].astype(np.float64).sum() / max(ref["v"][rng_].astype(np.float64).sum(), 1e-300)),
                "steps_mean_ours": float(ours["steps"][rng_].mean()), "steps_mean_ref": float(ref["steps"][rng_].mean()),
                "steps_differ": int(np.count_nonzero(ours["steps"][rng_] != ref["steps"][rng_])), "n": int(rng_.stop - rng_.start),
            }
            rep[name] = e
        report["levels"][str(s)] = rep
        print(f"--- step {s}: loss ours {losses[i].item():.6e} ref {ref_losses[i]:.6e}", flush=True)
        for name, e in rep.items():
            print(f"  {name:4s} n={e['n']:8d} g_rel={e['grad_rel_l2']:.3e} g_ratio={e['grad_abs_ratio']:.4f} cos={e['grad_cos']:.5f} nz o/r={e['grad_nz_ours']}/{e['grad_nz_ref']}"
                  f" only_o/r={e['grad_nz_only_ours']}/{e['grad_nz_only_ref']} d_rel={e['delta_rel_l2']:.3e} d_ratio={e['delta_norm_ratio']:.4f}"
                  f" m={e['m_rel_l2']:.2e} v={e['v_rel_l2']:.2e} vsum={e['v_sum_ratio']:.4f} steps o/r={e['steps_mean_ours']:.3f}/{e['steps_mean_ref']:.3f} diff={e['steps_differ']}", flush=True)
    ours_l = losses.cpu().numpy().astype(np.float64)
    report["loss_ratio_windows"] = {f"{lo}-{hi}": float(ours_l[lo:hi].mean() / ref_losses[lo:hi].astype(np.float64).mean())
                                    for lo, hi in ((0, 10), (10, 50), (50, 100), (100, 200), (200, 500), (500, 950), (950, 1000)) if hi <= steps}
    report["ref_vs_golden_last50"] = float(ref_losses[-50:].mean() / g["losses"][-50:].astype(np.float64).mean()) if steps == int(g["steps"]) else None
    print("loss ratio windows", report["loss_ratio_windows"], "ref run vs golden run (last 50)", report["ref_vs_golden_last50"], flush=True)
    golden["losses"] = ref_losses[:100]
    np.savez_compressed(os.path.join(outdir, "state_cfg1_2d_T19.npz"), **golden)
    np.save(os.path.join(outdir, "state_losses_ours.npy"), ours_l.astype(np.float32)); np.save(os.path.join(outdir, "state_losses_ref.npy"), ref_losses)
    json.dump(report, open(os.path.join(outdir, "state_compare.json"), "w"), indent=1)


def curve(outdir, tag):
    import torch
    import tinycudann as tcnn
    g = np.load(os.path.join(ROOT, "tests", "golden", "curve_cfg1_2d_T19.npz"))
    cfg = json.loads(str(g["config"]))
    B, n_in, n_out, steps, seed = (int(g[k]) for k in ("B", "n_in", "n_out", "steps", "seed"))
    X, T = batches(B, n_in, n_out, steps, seed, "/tmp")
    dev = torch.device("cuda:0")
    tr = tcnn.create_from_config(n_in, n_out, cfg, seed=1337, device=dev).trainer
    if os.environ.get("PROBE_UNFUSED"):
        tr.use_fused_step = False
    losses = torch.zeros(steps, dtype=torch.float32, device=dev)
    for i in range(steps):
        x, t = torch.from_numpy(np.ascontiguousarray(X[i])).to(dev), torch.from_numpy(np.ascontiguousarray(T[i])).to(dev)
        tr.training_step(x, t)
        losses[i] = tr.loss_value[0]
    ours = losses.cpu().numpy().astype(np.float64)
    ref = g["losses"].astype(np.float64)
    out = {"tag": tag, "last50_ratio": float(ours[-50:].mean() / ref[-50:].mean()),
           "windows": {f"{lo}-{hi}": float(ours[lo:hi].mean() / ref[lo:hi].mean()) for lo, hi in ((0, 10), (10, 50), (50, 100), (100, 200), (200, 500), (500, 950))}}
    print(json.dumps(out), flush=True)
    np.save(os.path.join(outdir, f"losses_{tag}.npy"), ours.astype(np.float32))
    with open(os.path.join(outdir, "curve_variants.jsonl"), "a") as f:
        f.write(json.dumps(out) + "\n")


if __name__ == "__main__":
    what, outdir = sys.argv[1], sys.argv[2]
    os.makedirs(outdir, exist_ok=True)
    if what == "flips":
        flips(outdir)
    elif what == "state":
        state(outdir)
    elif what == "curve":
        curve(outdir, sys.argv[3])
