"""One train step of each kernel family (for compute-sanitizer): base model batch 16 (one CTA per SM shapes), batch 304
(two-CTAs-per-SM shapes, TMA strips), fp32 mode, and the wide model batch 8."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
rng = np.random.default_rng(0)
for B, mode, dims in ((16, "bf16", {}), (304, "bf16", {}), (8, "fp32", {}), (8, "bf16", dict(c1=256, c2=256, f1=768, f2=384))):
    e = b200cnn.TrainStepEngine(batch=B, mode=mode, **dims); e.init_params(0)
    img = torch.as_tensor(rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8)).cuda()
    lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int32)).cuda()
    e.train_step(img, lab, 0.1)
    torch.cuda.synchronize()
    print(B, mode, dims, "loss", float(e.loss()), flush=True)
    del e
