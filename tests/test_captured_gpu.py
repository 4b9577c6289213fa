"""CapturedTrainStep: the CUDA-graph replay of the train step must follow the eager step exactly."""
import copy

import pytest
import torch

import pinot_b200 as pb
from pinot_b200 import synthetic
from conftest import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _net(seed=0):
    torch.manual_seed(seed)
    rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn(), [64, "tanh", 64, "tanh"])
    net = pb.Net(rep, output_regressor_class=pb.regressors.ExactGaussianProcessRegressor).to(DEV)
    with torch.no_grad():
        net.output_regressor.kernel.scale.fill_(-3.0)
    return net


def test_replay_follows_eager_training_and_load_swaps_the_batch():
    g, y = synthetic.make_batch(256, seed=11)
    g, y = g.to(DEV), y.to(DEV)
    eager = _net()
    captured = copy.deepcopy(eager)
    opt_e = torch.optim.Adam(eager.parameters(), 1e-3)
    opt_c = torch.optim.Adam(captured.parameters(), 1e-3, capturable=True)
    before = [p.detach().clone() for p in captured.parameters()]
    step = pb.CapturedTrainStep(captured, opt_c, g, y)
    for p, b in zip(captured.parameters(), before):
        assert torch.equal(p, b), "construction must not train"
    # a second batch with the same structure but different features / targets
    g2 = g.to(DEV)
    g2.ndata["h"] = g.ndata["h"] * 0.5 + 0.1
    g2.csr()
    y2 = -y
    for it in range(4):
        if it == 2:
            step.load(g2, y2)
        ge, ye = (g, y) if it < 2 else (g2, y2)
        opt_e.zero_grad(set_to_none=True)
        le = eager.loss(ge, ye).sum()
        le.backward()
        opt_e.step()
        lc = step()
        captured.output_regressor.verify()
        assert rel_err(lc, le) < 1e-5, it
    for (name, pe), (_, pc) in zip(eager.named_parameters(), captured.named_parameters()):
        assert rel_err(pc, pe) < 1e-4, name


def test_load_rejects_a_batch_of_another_size():
    g, y = synthetic.make_batch(64, seed=1)
    g, y = g.to(DEV), y.to(DEV)
    net = _net()
    step = pb.CapturedTrainStep(net, torch.optim.Adam(net.parameters(), 1e-3, capturable=True), g, y)
    other, yo = synthetic.make_batch(65, seed=2)
    with pytest.raises(ValueError, match="differ in size"):
        step.load(other.to(DEV), yo.to(DEV))


def test_captured_upload_replays_follow_the_pinned_host_buffers():
    g, y = synthetic.make_batch(200, seed=4)
    s, d = (t.clone().pin_memory() for t in g.edges())
    feat = g.ndata["h"].clone().pin_memory()
    yh = y.clone().pin_memory()
    stream = torch.cuda.Stream()
    up = pb.CapturedBatchUpload(s, d, feat, yh, g.number_of_nodes(), g.batch_num_nodes().tolist(),
                                g.batch_num_edges().tolist(), stream)
    for scale in (1.0, -2.0):
        feat.copy_(g.ndata["h"] * scale)          # the loader refills the pinned buffers in place
        yh.copy_(y * scale)
        gd, yd, ev = up.replay()
        ev.synchronize()
        want = g.to(DEV)
        want.ndata["h"] = want.ndata["h"] * scale
        assert torch.equal(yd, (y * scale).to(DEV))
        assert torch.equal(gd.ndata["h"], want.ndata["h"])
        for name in ("indptr", "indices", "indptr_t", "indices_t", "seg_offsets", "node_seg", "in_norm", "inv_in_deg"):
            assert torch.equal(getattr(gd.csr(), name), getattr(want.csr(), name)), name
    up.release()


def test_variational_gp_step_can_be_captured_and_follows_eager_training():
    """The variational head's step holds no device->host read any more (the strictly-lower scale is filled through
    explicit indices instead of masked_scatter): it captures, and the replay trains like the eager step."""
    torch.manual_seed(3)
    g, y = synthetic.make_batch(96, seed=21)
    g, y = g.to(DEV), y.to(DEV)
    rep = pb.representation.Sequential(pb.representation.dgl_legacy.gn("SAGEConv"), [64, "tanh", 64, "tanh"])
    eager = pb.Net(rep, output_regressor_class=pb.regressors.VariationalGaussianProcessRegressor, n_inducing_points=64).to(DEV)
    with torch.no_grad():
        eager.output_regressor.kernel.scale.fill_(-2.0)
        eager.output_regressor.y_tr_sigma_tril.normal_(0.0, 0.02)
    captured = copy.deepcopy(eager)
    # the explicit indices reproduce masked_scatter's layout
    head = eager.output_regressor
    m = head.y_tr_sigma_diag.shape[0]
    idx = torch.arange(m, device=DEV)
    want = torch.diag_embed(torch.exp(head.y_tr_sigma_diag)).masked_scatter(idx[:, None] > idx[None, :], head.y_tr_sigma_tril)
    assert torch.equal(head._y_tr_sigma(), want)
    opt_e = torch.optim.Adam(eager.parameters(), 1e-3)
    opt_c = torch.optim.Adam(captured.parameters(), 1e-3, capturable=True)
    step = pb.CapturedTrainStep(captured, opt_c, g, y)
    for it in range(3):
        opt_e.zero_grad()
        le = eager.loss(g, y).sum()
        le.backward()
        opt_e.step()
        lc = step()
        assert abs(float(lc) - float(le)) <= 1e-4 * max(1.0, abs(float(le))), (it, float(lc), float(le))
    captured.output_regressor.verify()
    for pe, pc in zip(eager.parameters(), captured.parameters()):
        assert rel_err(pc, pe) < 1e-4
    step.release()
