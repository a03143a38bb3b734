"""Three eager bf16 train steps at batch 128 (for ncu kernel captures)."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
import b200cnn
B = 128
e = b200cnn.TrainStepEngine(batch=B, mode="bf16"); e.init_params(0)
rng = np.random.default_rng(0)
img = torch.as_tensor(rng.integers(0, 256, (B, 3, 32, 32), dtype=np.uint8)).cuda()
lab = torch.as_tensor(rng.integers(0, 10, (B,)).astype(np.int32)).cuda()
for _ in range(3):
    e.train_step(img, lab, 0.1)
    torch.cuda.synchronize()
print("loss", float(e.loss()))
