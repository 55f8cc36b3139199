"""``Hyperband`` tuner (``kerastuner.tuners.Hyperband``; KerasTuner is the engine the reference's
``CloudTuner`` is built on, tuner/tuner.py:21-25, and this package re-creates that engine).

Successive halving in brackets: bracket ``s`` starts ``size(s, 0)`` random configurations on a small
epoch budget, keeps the best ``1 / factor`` of every round and trains them further - from the
checkpoint of the round before, not from scratch - until the last round runs ``max_epochs``.  Brackets
with fewer, longer-trained configurations follow; ``hyperband_iterations`` repeats the whole cycle.

The oracle tags every trial through reserved hyperparameter values:
``tuner/epochs``, ``tuner/initial_epoch``, ``tuner/bracket``, ``tuner/round`` and, for a promoted
configuration, ``tuner/trial_id`` (the trial it continues).  Several tuner processes can work on one
round at a time; a worker that finds nothing to do until the round is finished gets an ``IDLE`` trial.
"""
from __future__ import annotations

import math
import os
import time

from . import trial as trial_lib
from .oracle import RandomSearchOracle
from .tuner import Tuner

RESERVED = ("tuner/epochs", "tuner/initial_epoch", "tuner/bracket", "tuner/round", "tuner/trial_id")


class HyperbandOracle(RandomSearchOracle):
    def __init__(self, objective, max_epochs, factor=3, hyperband_iterations=1, seed=None,
                 hyperparameters=None, allow_new_entries=True, tune_new_entries=True):
        if factor < 2:
            raise ValueError("factor needs to be an int larger than 1.")
        super().__init__(objective, max_trials=None, hyperparameters=hyperparameters,
                         allow_new_entries=allow_new_entries, tune_new_entries=tune_new_entries, seed=seed)
        self.max_epochs = int(max_epochs)
        self.min_epochs = 1
        self.factor = int(factor)
        self.hyperband_iterations = int(hyperband_iterations or 1)
        self._iteration = 0
        self._bracket = self.num_brackets() - 1          # brackets run from the widest to the narrowest
        self._rounds = self._empty_rounds(self._bracket)

    # -- the schedule -------------------------------------------------------------------------------
    def num_brackets(self) -> int:
        epochs, n = self.max_epochs, 0
        while epochs >= self.min_epochs:
            epochs /= self.factor
            n += 1
        return n

    def num_rounds(self, bracket: int) -> int:
        return bracket + 1

    def round_size(self, bracket: int, round_num: int) -> int:
        """Configurations in round ``round_num`` of ``bracket``: every bracket ends with about the
        same number of survivors, earlier rounds hold ``factor`` times more each."""
        last = math.ceil(1 + math.log(self.max_epochs, self.factor)) / (bracket + 1)
        return int(math.ceil(last * self.factor ** (bracket - round_num)))

    def round_epochs(self, bracket: int, round_num: int) -> int:
        return int(math.ceil(self.max_epochs / self.factor ** (bracket - round_num)))

    def total_trials(self) -> int:
        one = sum(self.round_size(b, r) for b in range(self.num_brackets()) for r in range(self.num_rounds(b)))
        return one * self.hyperband_iterations

    def _empty_rounds(self, bracket):
        return [[] for _ in range(self.num_rounds(bracket))]     # per round: [{"id", "past_id"}]

    # -- trial creation -----------------------------------------------------------------------------
    def _score_of(self, trial_id):
        t = self.trials.get(trial_id)
        return None if t is None or t.status != trial_lib.TrialStatus.COMPLETED else t.score

    def _ranked(self, round_num):
        """Trial ids of a round, best first (failed trials drop out)."""
        reverse = self._objective().direction == "max"
        scored = [(self._score_of(e["id"]), e["id"]) for e in self._rounds[round_num]]
        return [i for _s, i in sorted((x for x in scored if x[0] is not None), key=lambda x: x[0],
                                      reverse=reverse)]

    def _capacity(self, round_num) -> int:
        """Planned size of the round, or the number of survivors of the round before if fewer."""
        size = self.round_size(self._bracket, round_num)
        if round_num == 0:
            return size
        return min(size, len(self._ranked(round_num - 1)))

    def _round_done(self, round_num) -> bool:
        if round_num > 0 and not self._round_done(round_num - 1):
            return False
        entries = self._rounds[round_num]
        if len(entries) < self._capacity(round_num):
            return False
        finished = (trial_lib.TrialStatus.COMPLETED, trial_lib.TrialStatus.INVALID)
        return all(self.trials[e["id"]].status in finished for e in entries)

    def _tagged(self, values, round_num, past_id=None):
        values = dict(values)
        values["tuner/epochs"] = self.round_epochs(self._bracket, round_num)
        values["tuner/initial_epoch"] = self.round_epochs(self._bracket, round_num - 1) if round_num else 0
        values["tuner/bracket"] = self._bracket
        values["tuner/round"] = round_num
        if past_id is not None:
            values["tuner/trial_id"] = past_id
        return values

    def _advance_bracket(self) -> bool:
        """Next bracket (or next iteration); False when the whole schedule is finished."""
        if self._bracket > 0:
            self._bracket -= 1
        else:
            self._iteration += 1
            if self._iteration >= self.hyperband_iterations:
                return False
            self._bracket = self.num_brackets() - 1
        self._rounds = self._empty_rounds(self._bracket)
        return True

    def _populate_space(self, trial_id):
        while True:
            rounds = self._rounds
            # 1. a fresh random configuration while the first round has room
            if len(rounds[0]) < self.round_size(self._bracket, 0):
                values = {hp.name: hp.default for hp in self.hyperparameters.space} \
                    if not self.trials else self._random_values()
                rounds[0].append({"id": trial_id, "past_id": None})
                return {"status": trial_lib.TrialStatus.RUNNING, "values": self._tagged(values, 0)}
            # 2. promote the best of a finished round into the next one
            for r in range(1, len(rounds)):
                if not self._round_done(r - 1):
                    break
                ranked = self._ranked(r - 1)
                room = self._capacity(r)
                if len(rounds[r]) >= room:
                    continue
                taken = {e["past_id"] for e in rounds[r]}
                for past_id in ranked[:room]:
                    if past_id not in taken:
                        values = {k: v for k, v in self.trials[past_id].hyperparameters.values.items()
                                  if k not in RESERVED}
                        rounds[r].append({"id": trial_id, "past_id": past_id})
                        return {"status": trial_lib.TrialStatus.RUNNING,
                                "values": self._tagged(values, r, past_id)}
            # 3. nothing to hand out: wait for running trials, or move on
            if any(t.status == trial_lib.TrialStatus.RUNNING for t in self.trials.values()):
                return {"status": trial_lib.TrialStatus.IDLE, "values": None}
            if not self._advance_bracket():
                return {"status": trial_lib.TrialStatus.STOPPED, "values": None}

    def _objective(self):
        return self.objective[0] if isinstance(self.objective, list) else self.objective

    # -- persistence ----------------------------------------------------------------------------------
    def get_state(self):
        state = super().get_state()
        state.update({"hyperband": {"iteration": self._iteration, "bracket": self._bracket,
                                    "rounds": self._rounds, "max_epochs": self.max_epochs,
                                    "factor": self.factor, "iterations": self.hyperband_iterations}})
        return state

    def set_state(self, state):
        super().set_state(state)
        hb = state.get("hyperband")
        if hb:
            self._iteration, self._bracket, self._rounds = hb["iteration"], hb["bracket"], hb["rounds"]


class Hyperband(Tuner):
    """``Hyperband(hypermodel, objective, max_epochs, factor=3, hyperband_iterations=1, ...)``.
    ``search(x, y, ...)`` takes the usual ``fit`` arguments; ``epochs`` and ``initial_epoch`` are set
    per trial by the schedule."""

    def __init__(self, hypermodel, objective, max_epochs, factor=3, hyperband_iterations=1, seed=None,
                 hyperparameters=None, tune_new_entries=True, allow_new_entries=True, **kwargs):
        oracle = HyperbandOracle(objective, max_epochs=max_epochs, factor=factor,
                                 hyperband_iterations=hyperband_iterations, seed=seed,
                                 hyperparameters=hyperparameters, tune_new_entries=tune_new_entries,
                                 allow_new_entries=allow_new_entries)
        super().__init__(oracle, hypermodel, **kwargs)

    def search(self, *fit_args, **fit_kwargs):
        self.on_search_begin()
        while True:
            trial = self.oracle.create_trial(self.tuner_id)
            if trial.status == trial_lib.TrialStatus.STOPPED:
                break
            if trial.status == trial_lib.TrialStatus.IDLE:
                time.sleep(0.05)                  # another worker still runs a trial of this round
                continue
            self.on_trial_begin(trial)
            self.run_trial(trial, *fit_args, **fit_kwargs)
            self.on_trial_end(trial)
        self.on_search_end()

    def run_trial(self, trial, *fit_args, **fit_kwargs):
        values = trial.hyperparameters.values
        fit_kwargs = dict(fit_kwargs)
        fit_kwargs["epochs"] = values["tuner/epochs"]
        fit_kwargs["initial_epoch"] = values["tuner/initial_epoch"]
        super().run_trial(trial, *fit_args, **fit_kwargs)

    def _build_model(self, hp):
        model = super()._build_model(hp)
        past_id = hp.values.get("tuner/trial_id")
        if past_id is not None:
            # continue the configuration's training: weights of the last epoch the parent ran
            base = os.path.join(self.get_trial_dir(past_id), "checkpoints")
            epochs = sorted(int(n.split("_", 1)[1]) for n in os.listdir(base) if n.startswith("epoch_")) \
                if os.path.isdir(base) else []
            if epochs and model.built:
                model.load_weights(os.path.join(base, "epoch_%d" % epochs[-1], "checkpoint"))
        return model
