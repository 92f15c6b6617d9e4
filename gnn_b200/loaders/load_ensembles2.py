"""Seeded-subsample ensemble datasets -- the surface of the reference's gnn/loaders/load_ensembles2.py.

``EnsembleGraph`` / ``load_data_ensembles2`` keep the reference's arguments and produce the same list of graphs
(``split`` training graphs drawn with ``df.sample(frac, random_state=split)``, each carrying ``.test``, plus an
optional "full" graph and a trailing validation graph).  The row draws are made on the host by pandas itself on a
frame that holds only row positions -- ``DataFrame.sample`` looks at nothing but the length -- so the indices are the
reference's bit for bit; the markers stay on the GPU and every graph is built by the same kernels as ``create_graph``.

Behaviour kept on purpose (all visible in the reference's output):
  * the validation rows are the first ``valid_size`` rows of the frame AFTER those rows were cut off (:82-83), so
    they overlap the training pool;
  * ``whole_train`` (:99) is drawn without a seed, from numpy's global generator (``np.random.seed`` controls it);
  * ``whole_test`` (:100) compares index LABELS with the re-numbered labels 0..m-1 of ``whole_train``: it is every
    row of the pool whose label is >= m, whether it was sampled for training or not;
  * ``hot=True`` raises the KeyError the reference raises at :139 (its loop runs past the marker count).
"""
import numpy as np

from .. import ops  # noqa: F401  (fails loudly when the CUDA library is missing)
from ..data import GenotypeTable
from .load import ProcessedDataset, check_metric, graph_from_markers


def _sample(n, frac, random_state=None):
    """Positions ``DataFrame.sample(frac=frac, random_state=random_state)`` picks in a frame of n rows."""
    import pandas as pd
    return pd.Series(np.arange(n)).sample(frac=frac, random_state=random_state).values


def ensemble_rows(n_rows, split, train_size=0.7, add_full=False, full_split=None, use_validation=False,
                  validation_size=0.1, separate_sets=True):
    """File rows of every graph of load_ensembles2.py:74-132: ([(train_rows, test_rows), ...], valid_rows | None).
    The pool keeps its original labels (= file positions) throughout, exactly like ``df_whole`` does."""
    pool = np.arange(n_rows)
    valid = None
    if use_validation:
        valid_size = validation_size if isinstance(validation_size, int) else int(n_rows * validation_size)
        pool = pool[valid_size:]
        valid = pool[:valid_size]
    if isinstance(full_split, float):
        full_split = int(full_split * len(pool))
    frac = train_size if train_size < 1 else train_size / len(pool)
    if separate_sets:
        whole_train = pool[_sample(len(pool), frac)]                # :99, unseeded
        whole_test = pool[pool >= len(whole_train)]                 # :100, by label
    out = []
    for s in range(split + (1 if add_full else 0)):
        if s == split and add_full:
            if full_split is None:
                train, test = pool[_sample(len(pool), frac, s)], pool
            else:
                train, test = pool[:full_split], pool[full_split:]
        elif separate_sets:
            train, test = whole_train[_sample(len(whole_train), 0.5, s)], whole_test
        elif full_split is None:
            train = pool[_sample(len(pool), frac, s)]
            test = pool[pool >= len(train)]
        else:
            less = pool[:full_split]
            frac = train_size if train_size < 1 else train_size / len(less)       # :123, stays in force afterwards
            train = less[_sample(len(less), frac, s)]
            test = less[less >= len(train)]
        out.append((train, test))
    return out, valid


class EnsembleGraph(ProcessedDataset):
    """load_ensembles2.py:15-297.  Same constructor arguments (plus ``device`` and ``cache``)."""

    def __init__(self, root, bits, raw_filename, num_neighbours=1, smoothing="laplacian", mode="connectivity",
                 scaled=True, remove_mean=True, split=5, hot=False, algorithm="euclidean", add_full=False,
                 full_split=None, train_size=0.7, use_weights=False, use_validation=False, validation_size=0.1,
                 include_self=False, separate_sets=True, device="cuda", cache=True):
        self.root = root
        self.num_neighbours = num_neighbours
        self.smoothing = smoothing
        self.mode = mode
        self.raw_filename = raw_filename
        self.bits = bits
        self.scaled = scaled
        self.remove_mean = remove_mean
        self.hot = hot
        self.algorithm = algorithm
        self.split = split
        self.train_size = train_size
        self.add_full = add_full
        self.full_split = full_split
        self.use_weights = use_weights
        self.use_validation = use_validation
        self.validation_size = validation_size
        self.include_self = include_self
        self.separate_sets = separate_sets
        self.device = device
        self._build(cache)

    @property
    def raw_file_names(self):
        return [self.raw_filename]

    @property
    def processed_file_names(self):
        """The cache key of load_ensembles2.py:44-48."""
        bits = "whole" if self.bits is None else "bits"
        filename = self.raw_filename.replace("/", ":")
        return [f"data-ensemble2-{filename}-{self.split}-{self.add_full}-{self.full_split}-{self.train_size}-{self.hot}-"
                f"{self.algorithm}-{self.num_neighbours}-{self.scaled}-{self.remove_mean}-{self.smoothing}-{self.mode}-"
                f"{self.use_weights}-{self.use_validation}-{self.include_self}-{self.separate_sets}-{bits}.pt"]

    def _graph(self, table):
        return graph_from_markers(table.markers, table.value, self.num_neighbours, self.smoothing, self.use_weights,
                                  self.scaled, self.include_self, self.mode, self.remove_mean)

    def process(self):
        import pandas as pd
        check_metric(self.algorithm, self.mode)
        data_list = []
        for filename in self.raw_file_names:
            with open(filename) as fp:
                column_count = len(fp.readline().split(","))
            labels = ["value"] + [str(i + 1) for i in range(column_count - 1)]
            df = pd.read_csv(filename, names=labels)               # :74-79
            if self.bits is not None:
                df = df[["value"] + self.bits]                     # :88-89
            if self.hot:
                raise KeyError(df.shape[1] - 1)                    # :139: conv has no entry past the last marker
            whole = GenotypeTable.from_frame(df, device=self.device)
            rows, valid = ensemble_rows(len(whole), self.split, self.train_size, self.add_full, self.full_split,
                                        self.use_validation, self.validation_size, self.separate_sets)
            tests = {}
            for train, test in rows:
                data = self._graph(whole.take(train))
                key = test.tobytes()
                if key not in tests:                               # whole_test is the same rows for every split
                    tests[key] = self._graph(whole.take(test))
                data.test = tests[key]
                assert data.edge_index.shape[0] > 0 and data.test.edge_index.shape[0] > 0
                data_list.append(data)
            if valid is not None:
                v = self._graph(whole.take(valid))
                v.test = v
                data_list.append(v)
        return data_list


def load_data_ensembles2(filename, bits=None, num_neighbours=1, smoothing="laplacian", mode="connectivity", scaled=True,
                         remove_mean=True, split=None, hot=False, algorithm="euclidean", add_full=False, train_size=0.7,
                         use_weights=False, use_validation=False, validation_size=0.1, full_split=None,
                         include_self=False, separate_sets=True, **kwargs):
    """load_ensembles2.py:300-302."""
    return EnsembleGraph(".", bits=bits, raw_filename=filename, num_neighbours=num_neighbours, smoothing=smoothing,
                         scaled=scaled, remove_mean=remove_mean, mode=mode, split=split, hot=hot, algorithm=algorithm,
                         add_full=add_full, train_size=train_size, use_weights=use_weights,
                         use_validation=use_validation, validation_size=validation_size, full_split=full_split,
                         include_self=include_self, separate_sets=separate_sets,
                         **{k: v for k, v in kwargs.items() if k in ("device", "cache")})


load_data = load_data_ensembles2
