"""Two estimator calls at (B, T) -- the short command profiled under ncu.  python tools/one_forward.py B T"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import serenade_b200 as S  # noqa: E402

B, T = int(sys.argv[1]), int(sys.argv[2])
dev = torch.device("cuda", 0)
cfm = S.CFM(in_channels=242, out_channels=80, spk_embed_dim=256, decoder_channels=(512, 512),
            decoder_attention_head_dim=512)
cfm.load_state_dict(S.make_state_dict(0, False), strict=True)
b = S.make_batch([T] * B, n_steps=2, seed=1235)
args = [b.z.to(dev), b.mask.to(dev), b.mu.to(dev), torch.tensor(0.3), b.spks.to(dev)]
for _ in range(2):
    v = cfm.estimator(*args)
torch.cuda.synchronize()
print("ok", float(v.abs().mean()))
