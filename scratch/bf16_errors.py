"""Measured bf16-mode errors against the fp64 oracle (profiles/r2_bf16_errors.json): max-abs-normalised relative error of
every forward tensor, the loss, and every gradient (gradients against the oracle's backward evaluated on the GPU's own
forward tensors, as in the tests), for raw-byte and well-conditioned inputs at batch 128 and 1024, base and wide model."""
import json, sys
import numpy as np, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import b200cnn
from oracle import ref_numpy as rn
from test_gpu_parity import _oracle_backward_on_gpu_forward

def relerr(got, want):
    want = np.asarray(want, np.float64)
    return float(np.abs(np.asarray(got, np.float64) - want).max() / (np.abs(want).max() + 1e-30))

out = {}
for model, dims in (("base", {}), ("wide", dict(c1=256, c2=256, f1=768, f2=384))):
    for kind in ("u8", "soft"):
        for B in ((128, 1024) if model == "base" else (128,)):
            rng = np.random.default_rng(41)
            img = rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8) if kind == "u8" else \
                np.clip(np.rint(rng.normal(8.0, 2.5, (B, 3, 32, 32))), 0, 255).astype(np.uint8)
            lab = rng.integers(0, 10, (B, 1)).astype(np.int32)
            params = rn.init_params(0, **dims)
            _, loss, g, a = rn.train_step(params, img, lab, lr=np.float32(0.1))
            e = b200cnn.TrainStepEngine(batch=B, mode="bf16", **dims)
            e.load_params(params)
            e.forward(torch.as_tensor(img).cuda(), torch.as_tensor(lab.reshape(-1)).cuda()); e.backward()
            torch.cuda.synchronize()
            r = {"forward": {n: relerr(e.tensor(n).cpu().numpy().reshape(a[n].shape), a[n])
                             for n in ("pool1", "pool2", "full1", "full2", "logits")},
                 "loss_rel": abs(float(e.loss()) - loss) / abs(loss)}
            g2, _ = _oracle_backward_on_gpu_forward(e, params, lab, a, fused_conv1=True)
            got = e.get_grads()
            r["grads_vs_oracle_on_gpu_forward"] = {k: relerr(got[k], g2[k]) for k in rn.PARAM_NAMES}
            r["grads_vs_pure_oracle"] = {k: relerr(got[k], g[k]) for k in rn.PARAM_NAMES}
            out[f"{model}_{kind}_b{B}"] = r
            print(model, kind, B, "fwd max %.2e loss %.2e grad max %.2e (pure oracle %.2e)" % (
                max(r["forward"].values()), r["loss_rel"], max(r["grads_vs_oracle_on_gpu_forward"].values()),
                max(r["grads_vs_pure_oracle"].values())), flush=True)
            del e
json.dump(out, open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/bf16_errors.json", "w"), indent=1)
