import sys
sys.path.insert(0, '/root/repo')
import torch
from seldon_core_b200 import ThompsonSamplingB200, synth, _cabi
T = 20000
p, u = synth.router_rewards(T, 32)
r = ThompsonSamplingB200(n_branches=32, seed=1)
dec = r.run_batched(T, torch.from_numpy(u).cuda(), reward_mode=_cabi.SB_REWARD_BERNOULLI, arm_p=torch.from_numpy(p).cuda())
torch.cuda.synchronize()
print(int(dec.sum()))
