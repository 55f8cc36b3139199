"""Metrics as (sum, count) accumulators: replicas add their local sums into the communicator's aux
scalars and the gradient all-reduce kernel sums them across replicas for free (K11; SUM-aggregated
metric variables of tf.distribute).  Reference usage:
``metrics=['accuracy']`` README.md:79, ``SparseCategoricalAccuracy`` testdata/mnist_example_using_ctl.py:63-66."""
from __future__ import annotations

import torch


class Metric:
    name = "metric"
    from_fused_xent = False    # True: the fused cross-entropy kernel already counts it
    extra_slots = 0            # > 0: the metric is built from 1 + extra_slots sums (``batch_vector`` / ``reduce_vector``)

    def __init__(self, name=None):
        if name:
            self.name = name
        self.reset_states()

    def batch_sum(self, y_true, y_pred) -> torch.Tensor:
        """Sum of the per-sample metric over the local batch (0-d tensor on the device)."""
        raise NotImplementedError

    # stand-alone use (custom training loops, mnist_example_using_ctl.py:63-66,139-141): the
    # accumulators are per replica - each replica sees its own shard of every batch, so result()
    # is that replica's estimate; sum (sum, count) pairs with strategy.reduce for a global number
    def reset_states(self):
        self._sum, self._count = 0.0, 0.0

    reset_state = reset_states

    def update_state(self, y_true, y_pred, sample_weight=None):
        y_true, y_pred = torch.as_tensor(y_true), torch.as_tensor(y_pred).detach()
        self._sum += float(self.batch_sum(y_true.to(y_pred.device), y_pred))
        self._count += float(y_pred.shape[0])

    def __call__(self, *a, **k):
        self.update_state(*a, **k)
        return self.result()

    def result(self):
        return self._sum / self._count if self._count else 0.0


class SparseCategoricalAccuracy(Metric):
    name = "sparse_categorical_accuracy"
    from_fused_xent = True

    def batch_sum(self, y_true, y_pred):
        return (y_pred.argmax(-1) == y_true.reshape(-1).to(y_pred.device)).sum().float()


class CategoricalAccuracy(Metric):
    name = "categorical_accuracy"

    def batch_sum(self, y_true, y_pred):
        return (y_pred.argmax(-1) == y_true.argmax(-1)).sum().float()


class BinaryAccuracy(Metric):
    name = "binary_accuracy"

    def batch_sum(self, y_true, y_pred):
        return ((y_pred.float() > 0.5).float() == y_true.float().reshape(y_pred.shape)).float() \
            .reshape(y_pred.shape[0], -1).mean(-1).sum()


class MeanAbsoluteError(Metric):
    name = "mae"

    def batch_sum(self, y_true, y_pred):
        d = (y_pred.float() - y_true.float().reshape(y_pred.shape)).abs()
        return d.reshape(d.shape[0], -1).mean(-1).sum()


class MeanSquaredError(Metric):
    name = "mse"

    def batch_sum(self, y_true, y_pred):
        d = (y_pred.float() - y_true.float().reshape(y_pred.shape)) ** 2
        return d.reshape(d.shape[0], -1).mean(-1).sum()


class SparseTopKCategoricalAccuracy(Metric):
    name = "sparse_top_k_categorical_accuracy"

    def __init__(self, k=5, name=None):
        self.k = int(k)
        super().__init__(name)

    def batch_sum(self, y_true, y_pred):
        top = y_pred.topk(min(self.k, y_pred.shape[-1]), dim=-1).indices
        return (top == y_true.reshape(-1, 1).to(top.device)).any(-1).sum().float()


class TopKCategoricalAccuracy(SparseTopKCategoricalAccuracy):
    name = "top_k_categorical_accuracy"

    def batch_sum(self, y_true, y_pred):
        return super().batch_sum(y_true.argmax(-1), y_pred)


class _RatioMetric(Metric):
    """Numerator and denominator are both sums over samples (and over replicas): precision = true
    positives / predicted positives, recall = true positives / actual positives."""
    extra_slots = 1

    def __init__(self, thresholds=0.5, class_id=None, name=None):
        self.threshold = 0.5 if thresholds is None else float(thresholds)
        self.class_id = class_id
        super().__init__(name)

    def _binary(self, y_true, y_pred):
        """(predicted positive, actually positive) as float masks of equal shape."""
        y_pred = y_pred.float()
        if y_pred.dim() > 1 and y_pred.shape[-1] > 1:            # multi-class scores
            if y_true.dim() == y_pred.dim() and y_true.shape[-1] == y_pred.shape[-1]:
                actual = y_true.float()                          # one-hot / multi-label targets
            else:
                actual = torch.nn.functional.one_hot(y_true.reshape(-1).long(), y_pred.shape[-1]).float()
            pred = (y_pred > self.threshold).float()
            if self.class_id is not None:
                pred, actual = pred[..., self.class_id], actual[..., self.class_id]
            return pred, actual.to(pred.device)
        return (y_pred.reshape(-1) > self.threshold).float(), y_true.float().reshape(-1).to(y_pred.device)

    def reset_states(self):
        self._sum, self._count = 0.0, 0.0

    reset_state = reset_states

    def update_state(self, y_true, y_pred, sample_weight=None):
        num, den = self.batch_sums(torch.as_tensor(y_true), torch.as_tensor(y_pred).detach())
        self._sum += float(num)
        self._count += float(den)

    def batch_sum(self, y_true, y_pred):
        return self.batch_sums(y_true, y_pred)[0]

    def batch_vector(self, y_true, y_pred):
        return torch.stack(self.batch_sums(y_true, y_pred))

    @staticmethod
    def reduce_vector(sums):
        return sums[0] / sums[1] if sums[1] > 0 else 0.0


class AUC(Metric):
    """Area under the ROC curve (``tf.keras.metrics.AUC``, ``curve="ROC"``, trapezoidal summation) for
    scores in [0, 1] and binary labels.  True and false positives are counted at ``num_thresholds``
    thresholds - sums over samples and replicas like every other metric - and the curve is integrated
    when the result is read.  At most 100 thresholds (the step's scalar vector has 256 entries)."""
    name = "auc"

    def __init__(self, num_thresholds=100, name=None, **kwargs):
        self.num_thresholds = max(3, min(int(num_thresholds), 100))
        self.extra_slots = 2 * self.num_thresholds + 1           # tp[T], fp[T], positives, negatives - 1
        eps = 1e-7
        inner = [(i + 1) / (self.num_thresholds - 1) for i in range(self.num_thresholds - 2)]
        self._thresholds = torch.tensor([0.0 - eps] + inner + [1.0 + eps])
        super().__init__(name)

    def batch_vector(self, y_true, y_pred):
        scores = y_pred.float().reshape(-1)
        actual = y_true.float().reshape(-1).to(scores.device)
        th = self._thresholds.to(scores.device)
        above = (scores[None, :] > th[:, None]).float()          # [T, batch]
        tp = above @ actual
        fp = above @ (1.0 - actual)
        return torch.cat([tp, fp, actual.sum()[None], (1.0 - actual).sum()[None]])

    def reduce_vector(self, sums):
        t = self.num_thresholds
        tp, fp, pos, neg = sums[:t], sums[t:2 * t], sums[2 * t], sums[2 * t + 1]
        if pos <= 0 or neg <= 0:
            return 0.0
        tpr = [v / pos for v in tp]
        fpr = [v / neg for v in fp]
        return float(sum((fpr[i] - fpr[i + 1]) * (tpr[i] + tpr[i + 1]) / 2.0 for i in range(t - 1)))

    def reset_states(self):
        self._vec = None

    reset_state = reset_states

    def update_state(self, y_true, y_pred, sample_weight=None):
        vec = self.batch_vector(torch.as_tensor(y_true), torch.as_tensor(y_pred).detach()).double()
        self._vec = vec if self._vec is None else self._vec + vec

    def batch_sum(self, y_true, y_pred):
        return self.batch_vector(y_true, y_pred)[0]

    def result(self):
        return 0.0 if self._vec is None else self.reduce_vector([float(v) for v in self._vec])


class Precision(_RatioMetric):
    name = "precision"

    def batch_sums(self, y_true, y_pred):
        pred, actual = self._binary(y_true, y_pred)
        return (pred * actual).sum(), pred.sum()


class Recall(_RatioMetric):
    name = "recall"

    def batch_sums(self, y_true, y_pred):
        pred, actual = self._binary(y_true, y_pred)
        return (pred * actual).sum(), actual.sum()


class Mean(Metric):
    """Running mean of whatever is fed (``tf.keras.metrics.Mean``)."""
    name = "mean"

    def update_state(self, values, sample_weight=None):
        v = torch.as_tensor(values).detach().float()
        self._sum += float(v.sum())
        self._count += float(v.numel())


def get(identifier, loss=None, output_dim=None):
    """Resolve Keras metric spellings; 'accuracy'/'acc' depends on the loss like in Keras."""
    if isinstance(identifier, Metric):
        return identifier
    if isinstance(identifier, type) and issubclass(identifier, Metric):
        return identifier()
    key = str(identifier).lower()
    if key in ("accuracy", "acc"):
        lname = getattr(loss, "name", "")
        if "sparse_categorical" in lname:
            m = SparseCategoricalAccuracy()
        elif "binary" in lname:
            m = BinaryAccuracy()
        elif "categorical" in lname:
            m = CategoricalAccuracy()
        else:
            m = SparseCategoricalAccuracy()
        m.name = "accuracy" if key == "accuracy" else "acc"
        return m
    table = {"sparse_categorical_accuracy": SparseCategoricalAccuracy,
             "categorical_accuracy": CategoricalAccuracy, "binary_accuracy": BinaryAccuracy,
             "mae": MeanAbsoluteError, "mean_absolute_error": MeanAbsoluteError,
             "mse": MeanSquaredError, "mean_squared_error": MeanSquaredError,
             "sparse_top_k_categorical_accuracy": SparseTopKCategoricalAccuracy,
             "top_k_categorical_accuracy": TopKCategoricalAccuracy,
             "precision": Precision, "recall": Recall, "auc": AUC}
    if key in table:
        return table[key]()
    raise ValueError(f"Unknown metric function: {identifier}")
