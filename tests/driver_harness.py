"""Runs the reference's OWN drivers -- ``networks.tester.render_test`` (tester.py:18-168) and
``networks.manipulator.manipulator_demo`` (manipulator.py:366-491) -- on a tiny synthetic scene, with the metrics and
image writers neutralised (oracle/ref_loader.functional_io_stubs), and dumps what the hot path returned per chunk.

    python tests/driver_harness.py {reference|installed} {cpu|cuda} OUT.npz

``reference``: the unmodified reference end to end (torch, any device) -- proves the harness drives the real code.
``installed``: after ``dm_nerf_b200.install()`` -- the same driver code, the B200 path underneath.
Run as a subprocess by tests/test_install.py (it patches ``sys.modules``; keep that out of the test process).
"""
import os
import sys
import tempfile
from types import SimpleNamespace

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

H, W, K_INS, N_TEST = 24, 32, 32, 256
NEAR, FAR, SC, NIMP = 4.0, 15.0, 64, 128


def main(which: str, device: str, out_path: str) -> None:
    import ref_loader
    from dm_nerf_b200 import synth            # seeded inputs only (pure torch-CPU code)
    root = ref_loader.reference_root()
    assert root is not None, "stage the reference with oracle/make_ref.sh"
    captured = ref_loader.functional_io_stubs()
    sys.path.insert(0, root)
    os.chdir(root)                             # the drivers open './data/color_dict.json'
    if which == "installed":
        import dm_nerf_b200
        assert dm_nerf_b200.install(reference_root=root) == "patched"
    dev = torch.device(device)
    # seeded inputs first, as numpy (the CPU generators must not see the CUDA default tensor type set below)
    sd_c, sd_f = synth.init_state_dict(K_INS, 1, 8.0), synth.init_state_dict(K_INS, 2, 8.0)
    K_np = synth.dmsr_camera(H, W, 0)[0].numpy()
    poses_np = np.stack([synth.dmsr_camera(H, W, v)[1].numpy() for v in range(2)], 0)
    gen = torch.Generator(device="cpu").manual_seed(5)
    imgs_np = torch.rand(2, H, W, 3, generator=gen, device="cpu").numpy()
    labels_np = synth.block_labels(H, W, 4, device="cpu")[None].repeat(2, 1, 1).numpy()
    if dev.type == "cuda":
        # exactly what config.initial() does on a CUDA machine (config.py:169-171)
        torch.set_default_tensor_type('torch.cuda.FloatTensor')

    # ---- the imports the drivers make (train_dmsr.py:5-11, test_dmsr.py:5-10), executed verbatim ----
    import config
    from networks.tester import render_test
    from networks import manipulator
    import networks.tester as tester_mod
    torch.autograd.set_detect_anomaly(False)

    args = SimpleNamespace(multires=10, multires_views=4, i_embed=0, netdepth=8, netwidth=256, ins_num=K_INS,
                           device=dev, datadir="./data/dmsr/bedroom", N_test=N_TEST, near=NEAR, far=FAR, N_samples=SC,
                           N_importance=NIMP, perturb=0.0, is_train=False, N_ins=None, target_label=None)
    pe, ve, mc, mf, args = config.create_nerf(args)
    mc.load_state_dict(sd_c)
    mf.load_state_dict(sd_f)
    mc, mf = mc.to(dev).eval(), mf.to(dev).eval()
    hwk = (H, W, K_np)
    # test_dmsr.py:23-25 builds its inputs with torch.Tensor(numpy): on the default device
    poses = torch.Tensor(poses_np)
    gt_imgs = torch.Tensor(imgs_np)
    gt_labels = torch.Tensor(labels_np).type(torch.int16)
    ins_rgbs = np.random.RandomState(0).randint(0, 255, size=(64, 3))

    # record what the hot path hands back to the driver, chunk by chunk
    rec = {"rgb": [], "ins": []}
    inner = tester_mod.dm_nerf

    def recording_dm_nerf(*a, **k):
        info = inner(*a, **k)
        rec["rgb"].append(info["rgb_fine"].detach().cpu())
        rec["ins"].append(info["ins_fine"].detach().cpu())
        return info
    tester_mod.dm_nerf = recording_dm_nerf
    if dev.type == "cuda":
        # the reference's AP code (evaluator.py:76-175, out of scope) was written for torch 1.8 and mixes CPU inputs
        # with default-device temporaries, which torch 2.x refuses; it is given the CPU default it needs for the call
        import networks.evaluator as eval_mod
        inner_eval = tester_mod.ins_eval

        def ins_eval_on_cpu(*a, **k):
            torch.set_default_tensor_type('torch.FloatTensor')
            keep, eval_mod.device = eval_mod.device, torch.device("cpu")          # evaluator.py:16 module global
            try:
                return inner_eval(*a, **k)
            finally:
                eval_mod.device = keep
                torch.set_default_tensor_type('torch.cuda.FloatTensor')
        tester_mod.ins_eval = ins_eval_on_cpu

    out = {}
    with tempfile.TemporaryDirectory() as tmp, torch.no_grad():
        render_test(pe, ve, mc, mf, poses, hwk, args, gt_imgs=gt_imgs, gt_labels=gt_labels, ins_rgbs=ins_rgbs,
                    savedir=tmp, matched_file=os.path.join(tmp, "matching_log.txt"))
        assert os.path.exists(os.path.join(tmp, "test_results.txt"))
        out["test_results"] = np.loadtxt(os.path.join(tmp, "test_results.txt"))
        out["render_rgb"] = torch.cat(rec["rgb"], 0).numpy()
        out["render_ins"] = torch.cat(rec["ins"], 0).numpy()
        out["render_png0"] = np.asarray(captured[os.path.join(tmp, "000.png")])
        out["n_chunks"] = np.array(len(rec["rgb"]))

        # ---- manipulator_demo: one view, one rigidly moved object (manipulator.py:430-433) ----
        th = 0.3
        trans = np.array([[np.cos(th), -np.sin(th), 0, 0.1], [np.sin(th), np.cos(th), 0, -0.2], [0, 0, 1, 0.05],
                          [0, 0, 0, 1]], dtype=np.float32)
        objs = [{"obj_name": "chair", "tar_id": 8, "mani_mode": "rotation"}]
        objs_trans = {"chair": [{"transformation": trans}]}
        captured.clear()
        torch.manual_seed(11)
        manipulator.manipulator_demo(pe, ve, mc, mf, None, hwk, objs_trans, tmp, ins_rgbs, objs, poses[:1], {}, args)
        out["mani_png0"] = np.asarray(captured[os.path.join(tmp, "mani_output", "0_rgb.png")])
        out["mani_mask0"] = np.asarray(captured[os.path.join(tmp, "mani_output", "0_ins_pred_mask.png")])
        out["mani_trans"] = trans
    out["bound"] = np.array([tester_mod.get_rays_k.__module__, manipulator.manipulator.__module__,
                             manipulator.manipulator_demo.__module__, type(mc).__module__])
    np.savez(out_path, **out)
    print("driver_harness ok:", which, device, {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3])
