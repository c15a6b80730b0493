"""Gradient-boosted decision trees as a trial: the non-neural family of the reference's example zoo
(``examples/decision_trees/gbt_titanic_estimator/model_def.py``: ``tf.estimator.BoostedTreesClassifier`` with ``n_trees``,
``max_depth``, ``learning_rate``, ``l1_regularization`` / ``l2_regularization``, ``min_node_weight`` hyperparameters).

There is no optimizer and no back-propagation, so this is not a ``PyTorchTrial``: :class:`BoostedTreesController` is a
``CallbackTrialController`` whose unit of work — one "batch" — is ONE BOOSTING ROUND over the whole training matrix. The
booster itself is written on torch tensor ops (it runs on the trial's GPU when there is one): features are quantised into
histogram bins once; every round computes the logistic loss's gradients / hessians, grows one oblivious tree level by level
— the split of a level is the (feature, bin) with the best summed second-order gain over the current leaves, found from
``scatter_add`` histograms — and adds the shrunken leaf values to the ensemble's margin.
"""
from __future__ import annotations

import pathlib
from typing import Any, Dict, List, Optional, Tuple

import torch

import determined_b200 as det


class ObliviousTree:
    """Depth-d tree with ONE (feature, threshold bin) per level: a sample's leaf is the d-bit number of its comparisons."""

    def __init__(self, features: torch.Tensor, bins: torch.Tensor, leaves: torch.Tensor) -> None:
        self.features, self.bins, self.leaves = features, bins, leaves

    def leaf_index(self, xb: torch.Tensor) -> torch.Tensor:
        idx = torch.zeros(xb.shape[0], dtype=torch.int64, device=xb.device)
        for f, b in zip(self.features.tolist(), self.bins.tolist()):
            idx = idx * 2 + (xb[:, f] > b).to(torch.int64)
        return idx

    def predict(self, xb: torch.Tensor) -> torch.Tensor:
        return self.leaves[self.leaf_index(xb)]


class GradientBoostedClassifier:
    """Binary classifier: margin = bias + sum of tree outputs, probability = sigmoid(margin)."""

    def __init__(self, max_depth: int = 6, learning_rate: float = 0.1, l1: float = 0.0, l2: float = 1.0, min_node_weight: float = 0.0,
                 n_bins: int = 32) -> None:
        self.max_depth, self.lr, self.l1, self.l2, self.min_node_weight, self.n_bins = max_depth, learning_rate, l1, l2, min_node_weight, n_bins
        self.edges: Optional[torch.Tensor] = None       # [F, n_bins - 1] bin boundaries (quantiles of the training matrix)
        self.bias = 0.0
        self.trees: List[ObliviousTree] = []

    # -- quantisation ---------------------------------------------------------------------------------------------------------
    def fit_bins(self, x: torch.Tensor) -> None:
        q = torch.linspace(0, 1, self.n_bins + 1, device=x.device)[1:-1]
        self.edges = torch.quantile(x.float(), q, dim=0).t().contiguous()

    def quantise(self, x: torch.Tensor) -> torch.Tensor:
        assert self.edges is not None, "fit_bins() first"
        return (x.float().unsqueeze(-1) > self.edges.to(x.device).unsqueeze(0)).sum(-1)        # [N, F] in [0, n_bins)

    # -- one boosting round -------------------------------------------------------------------------------------------------
    def _leaf_value(self, g: torch.Tensor, h: torch.Tensor) -> torch.Tensor:
        shrunk = torch.sign(g) * (g.abs() - self.l1).clamp_min(0.0)          # L1 soft-threshold on the gradient sum
        return -shrunk / (h + self.l2)

    def _score(self, g: torch.Tensor, h: torch.Tensor) -> torch.Tensor:
        shrunk = (g.abs() - self.l1).clamp_min(0.0)
        return shrunk * shrunk / (h + self.l2)

    def boost(self, xb: torch.Tensor, y: torch.Tensor, margin: torch.Tensor) -> Tuple[ObliviousTree, torch.Tensor]:
        """Grow one tree on the current gradients; returns it and the updated margin."""
        n, nf = xb.shape
        p = torch.sigmoid(margin)
        g, h = p - y, (p * (1 - p)).clamp_min(1e-6)
        leaf = torch.zeros(n, dtype=torch.int64, device=xb.device)
        feats, bins = [], []
        for depth in range(self.max_depth):
            n_leaf = 1 << depth
            # histograms of gradient / hessian sums per (leaf, feature, bin)
            flat = (leaf.unsqueeze(1) * nf + torch.arange(nf, device=xb.device)) * self.n_bins + xb
            hist_g = torch.zeros(n_leaf * nf * self.n_bins, device=xb.device).scatter_add_(0, flat.reshape(-1), g.unsqueeze(1).expand(n, nf).reshape(-1))
            hist_h = torch.zeros(n_leaf * nf * self.n_bins, device=xb.device).scatter_add_(0, flat.reshape(-1), h.unsqueeze(1).expand(n, nf).reshape(-1))
            hist_g, hist_h = hist_g.view(n_leaf, nf, self.n_bins), hist_h.view(n_leaf, nf, self.n_bins)
            gl, hl = hist_g.cumsum(-1), hist_h.cumsum(-1)                     # "bin <= b" goes left
            gt, ht = gl[..., -1:], hl[..., -1:]
            gain = self._score(gl, hl) + self._score(gt - gl, ht - hl) - self._score(gt, ht)
            gain = torch.where((hl >= self.min_node_weight) & (ht - hl >= self.min_node_weight), gain, torch.full_like(gain, -1.0))
            total = gain.clamp_min(0.0).sum(0)[:, :-1]                        # the last bin sends nothing right: not a split
            best = int(total.argmax())
            f, b = best // (self.n_bins - 1), best % (self.n_bins - 1)
            feats.append(f)
            bins.append(b)
            leaf = leaf * 2 + (xb[:, f] > b).to(torch.int64)
        n_leaf = 1 << self.max_depth
        sum_g = torch.zeros(n_leaf, device=xb.device).scatter_add_(0, leaf, g)
        sum_h = torch.zeros(n_leaf, device=xb.device).scatter_add_(0, leaf, h)
        values = self.lr * self._leaf_value(sum_g, sum_h)
        values = torch.where(sum_h > 0, values, torch.zeros_like(values))
        tree = ObliviousTree(torch.tensor(feats), torch.tensor(bins), values)
        self.trees.append(tree)
        return tree, margin + values[leaf]

    def margin(self, xb: torch.Tensor) -> torch.Tensor:
        m = torch.full((xb.shape[0],), self.bias, device=xb.device)
        for t in self.trees:
            m = m + t.leaves.to(xb.device)[t.leaf_index(xb)]
        return m

    # -- persistence --------------------------------------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, Any]:
        return {"edges": None if self.edges is None else self.edges.cpu(), "bias": self.bias,
                "trees": [(t.features, t.bins, t.leaves.cpu()) for t in self.trees]}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        self.edges, self.bias = sd["edges"], float(sd["bias"])
        self.trees = [ObliviousTree(f, b, l) for f, b, l in sd["trees"]]


def synthetic_titanic(size: int, seed: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """Passenger rows [sex, age, pclass, sibsp, parch, fare, embarked, alone] with survival drawn from the well-known pattern
    (women, children and first class first) plus noise."""
    g = torch.Generator().manual_seed(seed)
    sex = (torch.rand(size, generator=g) < 0.35).float()                      # 1 = female
    age = (29 + 14 * torch.randn(size, generator=g)).clamp(0.5, 80)
    pclass = torch.multinomial(torch.tensor([0.24, 0.21, 0.55]), size, True, generator=g).float() + 1
    sibsp = torch.poisson(torch.full((size,), 0.5), generator=g)
    parch = torch.poisson(torch.full((size,), 0.4), generator=g)
    fare = torch.exp(4.2 - 0.9 * pclass + 0.5 * torch.randn(size, generator=g))
    embarked = torch.multinomial(torch.tensor([0.72, 0.19, 0.09]), size, True, generator=g).float()
    alone = ((sibsp + parch) == 0).float()
    logit = -1.3 + 2.6 * sex - 0.9 * (pclass - 2) + 1.2 * (age < 12).float() - 0.015 * (age - 30) - 0.3 * (sibsp > 2).float()
    y = (torch.rand(size, generator=g) < torch.sigmoid(logit)).float()
    return torch.stack([sex, age, pclass, sibsp, parch, fare, embarked, alone], 1), y


class BoostedTreesController(det.CallbackTrialController):
    CHECKPOINT_FILENAME = "boosted_trees.pt"

    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(*args, **kwargs)
        hp = self.env.hparams
        self.device = torch.device("cuda", 0) if torch.cuda.is_available() and self.env.use_gpu else torch.device("cpu")
        self.model = GradientBoostedClassifier(int(hp.get("max_depth", 6)), float(hp.get("learning_rate", 0.1)),
                                               float(hp.get("l1_regularization", 0.0)), float(hp.get("l2_regularization", 1.0)),
                                               float(hp.get("min_node_weight", 0.0)), int(hp.get("n_bins", 32)))
        self.n_trees = int(hp.get("n_trees", 100))
        data = self.env.experiment_config.get("data") or {}
        xt, yt = synthetic_titanic(int(data.get("train_size", 891)), 1)
        xv, yv = synthetic_titanic(int(data.get("validation_size", 264)), 2)
        self.model.fit_bins(xt.to(self.device))
        self.xt, self.yt = self.model.quantise(xt.to(self.device)), yt.to(self.device)
        self.xv, self.yv = self.model.quantise(xv.to(self.device)), yv.to(self.device)
        prior = float(yt.mean().clamp(1e-3, 1 - 1e-3))
        self.model.bias = float(torch.log(torch.tensor(prior / (1 - prior))))
        if self.load_path is not None:
            self.load(pathlib.Path(self.load_path))
        self.margin = self.model.margin(self.xt)

    @classmethod
    def from_trial(cls, trial_inst: Any, context: Any, env: Any, workloads: Any, load_path: Any = None, rendezvous_info: Any = None,
                   **_: Any) -> "BoostedTreesController":
        return cls(context, env, workloads, load_path, rendezvous_info)

    @staticmethod
    def pre_execute_hook(env: Any, dist: Any = None) -> None:
        torch.manual_seed(env.trial_seed % (2 ** 31))

    @staticmethod
    def _metrics(margin: torch.Tensor, y: torch.Tensor) -> Dict[str, float]:
        loss = torch.nn.functional.binary_cross_entropy_with_logits(margin, y)
        return {"loss": float(loss), "accuracy": float(((margin > 0).float() == y).float().mean())}

    def train_for_step(self, step_id: int, num_batches: int) -> Dict[str, Any]:
        per_round = []
        for _ in range(num_batches):
            if len(self.model.trees) < self.n_trees:          # once the ensemble is full, further rounds change nothing
                _, self.margin = self.model.boost(self.xt, self.yt, self.margin)
            per_round.append(self._metrics(self.margin, self.yt))
        return {"metrics": det.util.make_metrics(int(self.xt.shape[0]) * num_batches, per_round),
                "stop_requested": self.context.get_stop_requested()}

    def compute_validation_metrics(self, step_id: int) -> Dict[str, Any]:
        m = self._metrics(self.model.margin(self.xv), self.yv)
        return {"metrics": {"validation_metrics": {"loss": m["loss"], "accuracy": m["accuracy"], "n_trees": len(self.model.trees)},
                            "num_inputs": int(self.xv.shape[0])},
                "stop_requested": self.context.get_stop_requested()}

    def save(self, path: pathlib.Path) -> None:
        path.mkdir(parents=True, exist_ok=True)
        torch.save(self.model.state_dict(), str(path / self.CHECKPOINT_FILENAME))

    def load(self, path: pathlib.Path) -> None:
        self.model.load_state_dict(torch.load(str(path / self.CHECKPOINT_FILENAME), weights_only=False))


class GBTTitanicTrial(det.Trial):
    trial_controller_class = BoostedTreesController

    def __init__(self, context: det.TrialContext) -> None:
        self.context = context
