"""Two eager bf16 train steps each of the batch-4096 base model and the batch-128 wide model (for ncu kernel captures)."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
rng = np.random.default_rng(0)
for B, dims in ((4096, {}), (128, dict(c1=256, c2=256, f1=768, f2=384))):
    e = b200cnn.TrainStepEngine(batch=B, mode="bf16", **dims); e.init_params(0)
    img = torch.as_tensor(rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8)).cuda()
    lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int32)).cuda()
    for _ in range(2):
        e.train_step(img, lab, 0.1)
        torch.cuda.synchronize()
    print(B, dims, "loss", float(e.loss()))
    del e
