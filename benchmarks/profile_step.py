"""A few EAGER training steps of the headline NCF config, for `ncu` launch lists / captures.
usage: python benchmarks/profile_step.py [steps] [scale]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from zoo.common.nncontext import init_nncontext
from zoo.feature.common import MiniBatch
from zoo.models.recommendation import NeuralCF
from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
from zoo.pipeline.estimator import Estimator

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
ctx = init_nncontext()
U, I, B = int(12_500_000 * scale), int(1_250_000 * scale), 65536
m = NeuralCF(U, I, 5, user_embed=128, item_embed=128, hidden_layers=[256, 128, 64], mf_embed=64, device=ctx.device)
est = Estimator.from_torch(model=m, optimizer=torch.optim.Adam(m.parameters(), lr=1e-3),
                           loss=SparseCategoricalCrossEntropy(zero_based_label=False, padding_value=None), ctx=ctx)
est.engine.use_cuda_graph = False
est.engine.prepare()
g = torch.Generator().manual_seed(0)
for s in range(steps):
    x = torch.stack([torch.randint(1, U + 1, (B,), generator=g), torch.randint(1, I + 1, (B,), generator=g)], 1).cuda()
    y = torch.randint(1, 6, (B,), generator=g).cuda()
    loss = est.engine.train_step(est.loss.cuda(), MiniBatch(x, y, B))
torch.cuda.synchronize()
print("loss", float(loss))
