"""CPU restatement of ``pinot.representation.Sequential`` + ``pinot.Net`` on the oracle's DGL
restatement and GP restatement.  ORACLE -- test infrastructure only.

Parameters are addressed by the reference's ``state_dict`` keys (SURVEY.md section 5:
``representation.f_in.0.{weight,bias}``, ``representation.d{idx}.gn.{weight,bias}`` or
``...gn.fc_{self,neigh}.{weight,bias}``, ``output_regressor.{log_sigma, kernel.scale, kernel.variance}``).
"""
import torch

from . import dgl_ref, gp_ref


def graph_from_arrays(src, dst, num_nodes, batch_num_nodes, batch_num_edges, feats):
    return dgl_ref.from_arrays(src, dst, num_nodes, batch_num_nodes, batch_num_edges, ndata={"h": feats})


def representation(g, params, config, layer="GraphConv", prefix="representation.", pool=True):
    """pinot/representation/sequential.py:78-97: f_in (Linear + tanh), then the config entries in
    order (int -> conv layer, str -> torch.<activation>), then sum-pooling over each graph's nodes."""
    x = g.ndata["h"].to(params[prefix + "f_in.0.weight"].dtype)
    x = torch.tanh(x @ params[prefix + "f_in.0.weight"].t() + params[prefix + "f_in.0.bias"])
    for idx, exe in enumerate(config):
        if isinstance(exe, int):
            key = prefix + "d%d.gn." % idx
            if layer == "GraphConv":
                x = dgl_ref.graph_conv(g, x, params[key + "weight"], params[key + "bias"], "both")
            elif layer == "SAGEConv":
                x = dgl_ref.sage_conv_mean(
                    g, x, params[key + "fc_self.weight"], params[key + "fc_self.bias"],
                    params[key + "fc_neigh.weight"], params[key + "fc_neigh.bias"])
            else:
                raise ValueError(layer)
        elif isinstance(exe, str):
            x = getattr(torch, exe)(x)
        else:
            raise ValueError("dropout entries are not part of the parity configurations")
    if not pool:
        return x
    with g.local_scope():
        g.ndata["h"] = x
        return dgl_ref.sum_nodes(g, "h")


def exact_net_loss(g, y, params, config, layer="GraphConv", kernel="rbf", logdet="trace"):
    """pinot/net.py:170-195 with an ExactGaussianProcessRegressor head."""
    h = representation(g, params, config, layer)
    return gp_ref.exact_gp_loss(
        h, y, params["output_regressor.kernel.scale"], params["output_regressor.kernel.variance"],
        params["output_regressor.log_sigma"], kernel=kernel, logdet=logdet), h


def exact_net_condition(g_te, g_tr, y_tr, params, config, layer="GraphConv", kernel="rbf"):
    """pinot/net.py:205-248: the training graphs are re-embedded and conditioned on."""
    h_te = representation(g_te, params, config, layer)
    h_tr = representation(g_tr, params, config, layer)
    return gp_ref.exact_gp_condition(
        h_te, h_tr, y_tr, params["output_regressor.kernel.scale"], params["output_regressor.kernel.variance"],
        params["output_regressor.log_sigma"], kernel=kernel)


def vgp_params(params, prefix="output_regressor."):
    keys = ["kernel.scale", "kernel.variance", "log_sigma", "x_tr", "y_tr_mu", "y_tr_sigma_diag", "y_tr_sigma_tril"]
    return {k: params[prefix + k] for k in keys}
