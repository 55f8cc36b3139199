"""``HyperModel``: something with ``build(hp) -> model`` (KerasTuner engine)."""
from __future__ import annotations


class HyperModel:
    def __init__(self, name=None, tunable=True):
        self.name = name
        self.tunable = tunable

    def build(self, hp):
        raise NotImplementedError


class DefaultHyperModel(HyperModel):
    def __init__(self, build, name=None, tunable=True):
        super().__init__(name=name, tunable=tunable)
        self.build = build


def get_hypermodel(hypermodel):
    if isinstance(hypermodel, HyperModel):
        return hypermodel
    if not callable(hypermodel):
        raise ValueError("The `hypermodel` argument should be either a callable with signature "
                         "`build(hp)` returning a model, or an instance of `HyperModel`.")
    return DefaultHyperModel(hypermodel)
