"""Config objects with the reference's interface (misc/utils.py:13-61): JSON -> attributes, `.dict` view.
Layer code injects defaults lazily by mutating `params.dict` (model/tdnn.py:114-116,152-154,163-164,173-174);
that behaviour is part of the observable API and is kept."""
import json


class Params():
    """Loads hyperparameters from a json file (reference: misc/utils.py:13-41)."""

    def __init__(self, json_path):
        self.update(json_path)

    def save(self, json_path):
        with open(json_path, 'w') as f:
            json.dump(self.__dict__, f, indent=4)

    def update(self, json_path):
        with open(json_path) as f:
            params = json.load(f)
            self.__dict__.update(params)

    @property
    def dict(self):
        return self.__dict__


class ParamsPlain():
    """Hyperparameters set by hand (reference: misc/utils.py:44-61)."""

    def __init__(self):
        pass

    @property
    def dict(self):
        return self.__dict__


def compute_cos_pairwise_eer(embeddings, labels, max_num_embeddings=1000):
    """misc/utils.py:307-346: equal error rate of the cosine scores of all pairs of (a subsample of at most
    max_num_embeddings of) the embeddings, target = same label; the metric train.py:119-123 logs after every validation
    pass (Trainer.valid(..., output_embeddings=True) supplies the arguments).  ROC by sklearn, the crossing point by
    interp1d + brentq as in the reference; the pair list is built with array indexing instead of a double Python loop, the
    caller's array is not normalised in place and no score file is written."""
    import numpy as np
    from scipy.interpolate import interp1d
    from scipy.optimize import brentq
    from sklearn import metrics
    emb = np.asarray(embeddings, np.float64)
    labels = np.asarray(labels)
    emb = emb / np.sqrt(np.sum(emb ** 2, axis=1, keepdims=True) + 1e-12)
    n = emb.shape[0]
    if n > max_num_embeddings:
        step = n // max_num_embeddings                      # Python-2 integer division in the reference
        emb, labels = emb[::step], labels[::step]
        n = emb.shape[0]
    i, j = np.triu_indices(n, k=1)
    scores = np.sum(emb[i] * emb[j], axis=1) if n < 64 else (emb @ emb.T)[i, j]
    keys = (labels[i] == labels[j]).astype(np.float64)
    fpr, tpr, _ = metrics.roc_curve(keys, scores, pos_label=1)
    return float(brentq(lambda x: 1.0 - x - interp1d(fpr, tpr)(x), 0.0, 1.0))
