"""Metric bookkeeping (host side; API mirror of the reference's iit/utils/metric.py:5-88)."""
from __future__ import annotations

from enum import Enum

import numpy as np


class MetricType(Enum):
    ACCURACY = 1
    LOSS = 2
    LOG = 3


class MetricStore:
    def __init__(self, name, metric_type: MetricType):
        if metric_type not in MetricType:
            raise AssertionError(f"Invalid metric type {metric_type}")
        self._name, self.type, self._store = name, metric_type, []

    def append(self, metric):
        self._store.append(metric)

    def get_value(self):
        if not self._store:
            raise ValueError("No values in metric store!")
        mean = np.mean(self._store)
        return mean * 100 if self.type == MetricType.ACCURACY else mean  # accuracies are mean-of-batch-means x100

    def get_name(self):
        return self._name

    def __str__(self) -> str:
        if self.type == MetricType.ACCURACY:
            return f"{self._name}: {float(self.get_value()):.2f}%"
        return f"{self._name}: {self.get_value():.4f}"

    def __len__(self):
        return len(self._store)


class PerTokenMetricStore(MetricStore):
    def __init__(self, name, **kwargs):
        super().__init__(name, metric_type=MetricType.LOG)
        np.set_printoptions(precision=kwargs.get("precision", 3))

    def get_value(self):
        if not self._store:
            raise ValueError("No values in metric store!")
        return np.mean(self._store, axis=0)

    def __str__(self) -> str:
        return f"{self._name}: {self.get_value()}"


class MetricStoreCollection:
    def __init__(self, list_of_metric_stores):
        self.metrics = list_of_metric_stores

    def update(self, metrics: dict):
        by_name = {m.get_name(): m for m in self.metrics}
        for k, v in metrics.items():
            if k not in by_name:
                raise AssertionError(f"Key {k} not found in metric stores!")
            by_name[k].append(v)
        lengths = {len(m) for m in self.metrics}
        if len(lengths) != 1:
            raise AssertionError(f"All metric stores should have the same length after update!, got lengths: "
                                 f"{[len(m) for m in self.metrics]}")

    def create_metric_store(self, name, metric_type: MetricType):
        store = MetricStore(name, metric_type)
        self.metrics.append(store)
        return store

    def __str__(self) -> str:
        return "\n".join(str(m) for m in self.metrics)

    def to_dict(self):
        return {m.get_name(): m.get_value() for m in self.metrics}
