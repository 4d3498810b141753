"""Loss closures and parameter groups: the caller side of the drop-in boundary, mirroring
`cmf/non_square_helpers.py:8-135` (same function names, config keys, asserts and return shapes)
so ``Trainer._train_batch`` (`cmf/trainer.py:191-224`) can use either implementation.
"""
from __future__ import annotations

import numpy as np


def get_non_square_parameters(module, m_flow):
    if m_flow:
        raise ValueError("m_flow (the ManifoldFlow baseline head) is outside the fused hot path")
    return [module.parameters()]


def get_non_square_train_metrics(config):
    # m_flow alternates two objectives per "epoch"; kept for the schedule arithmetic only
    num_objectives = 2 if config["m_flow"] else 1
    if config["likelihood_warmup"]:
        warmup_bounds = [num_objectives * config["likelihood_warmup_start"],
                         num_objectives * config["likelihood_warmup_end"]]
        likelihood_introduction_epoch, early_stopping_start_epoch = warmup_bounds
    else:
        warmup_bounds = None
        likelihood_introduction_epoch = early_stopping_start_epoch = 0

    def schedule(epoch):
        if config["likelihood_warmup"]:
            lw = np.interp(epoch, warmup_bounds, [0, 1]) if (epoch + 1) % num_objectives == 0 else 0
        else:
            lw = float((epoch + 1) % num_objectives == 0)
        return lw, epoch % num_objectives == 0

    def train_metrics(density, x, epoch):
        lw, add_rec = schedule(epoch)
        loss = -density.elbo(x, likelihood_wt=lw, add_reconstruction=add_rec)["elbo"].mean()
        return {"loss": loss}

    def train_metrics_l1_diagonal(density, x, epoch):
        lw, add_rec = schedule(epoch)
        loss = -density.elbo(
            x,
            likelihood_wt=lw * config["elbo_regularization_param"],
            metric_wt=lw * config["metric_regularization_param"],
            add_reconstruction=add_rec,
            add_diagonal_metric_reg=add_rec,
            add_offdiagonal_metric_reg=False,
        )["elbo"].mean()
        return {"loss": loss}

    def train_metrics_l1_offdiagonal(density, x, epoch):
        lw, add_rec = schedule(epoch)
        loss = -density.elbo(
            x,
            likelihood_wt=lw * config["elbo_regularization_param"],
            metric_wt=lw * config["metric_regularization_param"],
            add_reconstruction=add_rec,
            add_diagonal_metric_reg=False,
            add_offdiagonal_metric_reg=add_rec,
        )["elbo"].mean()
        return {"loss": loss}

    if config["g_kk_loss"]:
        assert config["g_ij_loss"] == False, "Cannot have both diagonal and offdiagonal terms in l1 yet.  Exiting..."
        return train_metrics_l1_diagonal, likelihood_introduction_epoch, early_stopping_start_epoch
    elif config["g_ij_loss"]:
        assert config["latent_dimension"] != 1, "There is no offdiagonal for 1d latent. Exiting..."
        return train_metrics_l1_offdiagonal, likelihood_introduction_epoch, early_stopping_start_epoch
    return train_metrics, likelihood_introduction_epoch, early_stopping_start_epoch
