"""Flagship model on the GPU kernels: parity with the CPU fp32 path and convergence."""
import copy

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _data(U, I, C, n, seed=0):
    rng = np.random.RandomState(seed)
    x = np.stack([rng.randint(1, U + 1, n), rng.randint(1, I + 1, n)], 1).astype(np.int64)
    y = ((x[:, 0] * 7 + x[:, 1] * 3) % C + 1).astype(np.int64)
    return x, y


@pytest.mark.parametrize("dims", [dict(user_embed=20, item_embed=20, hidden_layers=[40, 20, 10], mf_embed=20),
                                  dict(user_embed=64, item_embed=64, hidden_layers=[128, 64], mf_embed=32)])
def test_forward_matches_cpu(dims):
    from zoo.models.recommendation import NeuralCF

    torch.manual_seed(0)
    m = NeuralCF(300, 200, 5, **dims)
    x, _ = _data(300, 200, 5, 1000)
    cpu = m(torch.from_numpy(x)).detach()
    g = copy.deepcopy(m).cuda()
    gpu = g(torch.from_numpy(x).cuda()).detach().cpu()
    torch.testing.assert_close(gpu, cpu, atol=2e-2, rtol=5e-2)


@pytest.mark.parametrize("sparse", [True, False])
def test_ncf_converges_on_gpu(sparse):
    from zoo.common.nncontext import init_nncontext
    from zoo.common.triggers import MaxEpoch
    from zoo.feature.common import FeatureSet
    from zoo.models.recommendation import NeuralCF
    from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
    from zoo.pipeline.api.keras.optimizers import Adam
    from zoo.pipeline.estimator import Estimator

    torch.manual_seed(0)
    ctx = init_nncontext()
    U, I, C = 200, 100, 5
    x, y = _data(U, I, C, 20000)
    model = NeuralCF(U, I, C, user_embed=32, item_embed=32, hidden_layers=[64, 32, 16], mf_embed=16, sparse=sparse)
    est = Estimator(model, Adam(lr=1e-2))
    fs = FeatureSet.numpy(x, y)
    est.train(fs, SparseCategoricalCrossEntropy(zero_based_label=False), MaxEpoch(12), batch_size=1000)
    res = est.evaluate(fs, ["accuracy"], batch_size=5000)
    assert res[0].value > 0.9, res
    assert est.engine.state["Loss"] < 0.5


def test_smoke_entry():
    import __graft_entry__ as g

    g.smoke()


def test_cuda_graph_step_matches_eager():
    from zoo.common.nncontext import init_nncontext
    from zoo.feature.common import MiniBatch
    from zoo.models.recommendation import NeuralCF
    from zoo.pipeline.api.keras.objectives import SparseCategoricalCrossEntropy
    from zoo.pipeline.estimator import Estimator

    ctx = init_nncontext()
    torch.manual_seed(0)
    base = NeuralCF(500, 300, 5, user_embed=64, item_embed=64, hidden_layers=[128, 64], mf_embed=32)
    x, y = _data(500, 300, 5, 8 * 1024)
    xs = torch.from_numpy(x).cuda().split(1024)
    ys = torch.from_numpy(y).cuda().split(1024)
    results = []
    for use_graph in (False, True):
        m = copy.deepcopy(base)
        # SGD: the update is linear in the gradient, so the fp32-atomic rounding noise of two runs stays at rounding size
        # (Adam would turn a sign flip of a ~0 gradient into a 2*lr step and make the comparison flaky)
        est = Estimator.from_torch(model=m, optimizer=torch.optim.SGD(m.parameters(), lr=0.1),
                                   loss=SparseCategoricalCrossEntropy(zero_based_label=False, padding_value=None), ctx=ctx)
        est.engine.use_cuda_graph = use_graph
        est.engine.prepare()
        losses = []
        for a, b in zip(xs, ys):
            losses.append(float(est.engine.train_step(est.loss.cuda(), MiniBatch(a, b, 1024))))
        torch.cuda.synchronize()
        assert (est.engine._graph is not None and est.engine._graph["graph"] is not None) == use_graph
        results.append((losses, m.linear_last.weight.detach().clone(), m.mlp_user_embed.weight.detach().clone()))
    (l0, w0, e0), (l1, w1, e1) = results
    torch.testing.assert_close(torch.tensor(l0), torch.tensor(l1), atol=2e-3, rtol=2e-3)
    # fp32 atomics in wgrad / duplicate accumulation make the two runs differ only in rounding
    torch.testing.assert_close(w0, w1, atol=2e-3, rtol=2e-2)
    torch.testing.assert_close(e0, e1, atol=2e-3, rtol=2e-2)
