"""experiment -- host-side mirror of the reference's code/experiment.py.

Experiment is file-system bookkeeping (directory, log, dill dumps) and stays plain Python, as in the
reference.  The two hot loops of this module -- FixpointExperiment.run_net (code/experiment.py:70-77) and
count (:79-91) -- are ONE kernel launch each here: the stop rule, the per-step trajectory and the 5-way
classification run on the device (srnn_self_apply mode=SRNN_RUN_UNTIL, srnn_classify).

Star-import surface kept for the setups: os, time, dill, tqdm, copy (code/experiment.py:1-5).
"""
import copy
import os
import time

import dill
from tqdm import tqdm

__all__ = ["os", "time", "dill", "tqdm", "copy", "Experiment", "FixpointExperiment", "MixedFixpointExperiment",
           "SoupExperiment", "IdentLearningExperiment"]


class Experiment:
    """Context manager that owns `experiments/exp-{name}-{id}-{it}/`, a log and dill dumps
    (code/experiment.py:8-59)."""

    @staticmethod
    def from_dill(path):
        with open(path, "rb") as fh:
            return dill.load(fh)

    def __init__(self, name=None, ident=None):
        self.experiment_id = "{}_{}".format(ident or "", time.time())
        self.experiment_name = name or "unnamed_experiment"
        self.next_iteration = 0
        self.log_messages = []
        self.historical_particles = {}

    def __enter__(self):
        leaf = "exp-{name}-{id}-{it}".format(name=self.experiment_name, id=self.experiment_id,
                                             it=self.next_iteration)
        self.dir = os.path.join("experiments", leaf)
        os.makedirs(self.dir)                       # must not pre-exist, as in the reference
        print("** created {dir} **".format(dir=self.dir))
        return self

    def __exit__(self, exc_type, exc_value, traceback):
        self.save(experiment=self.without_particles())
        self.save_log()
        self.next_iteration += 1

    def log(self, message, **kwargs):
        self.log_messages.append(message)
        print(message, **kwargs)

    def save_log(self, log_name="log"):
        path = os.path.join(self.dir, "{name}.txt".format(name=log_name))
        with open(path, "w") as fh:
            for message in self.log_messages:
                print(str(message), file=fh)

    def __copy__(self):
        twin = Experiment(name=self.experiment_name)
        twin.__dict__ = {k: v for k, v in self.__dict__.items() if k not in ("particles", "historical_particles")}
        return twin

    def without_particles(self):
        """Picklable copy: particles replaced by their `states` lists (code/experiment.py:50-54)."""
        twin = copy.copy(self)
        twin.historical_particles = {uid: p.states for uid, p in self.historical_particles.items()}
        return twin

    def save(self, **kwargs):
        for name, value in kwargs.items():
            with open(os.path.join(self.dir, "{name}.dill".format(name=name)), "wb") as fh:
                dill.dump(value, fh)


class FixpointExperiment(Experiment):

    def __init__(self, **kwargs):
        kwargs.setdefault("name", self.__class__.__name__)
        super().__init__(**kwargs)
        self.counters = dict(divergent=0, fix_zero=0, fix_other=0, fix_sec=0, other=0)
        self.interesting_fixpoints = []

    def run_net(self, net, step_limit=100, run_id=0):
        """while i < step_limit and not diverged and not fixpoint: self_attack [; save_state(time=i)]; then count.
        One fused launch when `net` is an engine net (always, unless a user subclass overrides the operator)."""
        fused = getattr(net, "_engine_run_until", None)
        if fused is not None and net._fused():
            fused(step_limit, log_states=bool(run_id))
        else:
            i = 0
            while i < step_limit and not net.is_diverged() and not net.is_fixpoint():
                net.self_attack()
                i += 1
                if run_id:
                    net.save_state(time=i)
        self.count(net)

    def count(self, net):
        label = getattr(net, "_engine_class_name", None)
        if label is not None and net._fused():
            label = label()
        elif net.is_diverged():
            label = "divergent"
        elif net.is_fixpoint():
            label = "fix_zero" if net.is_zero() else "fix_other"
        elif net.is_fixpoint(2):
            label = "fix_sec"
        else:
            label = "other"
        self.counters[label] += 1
        if label == "fix_other":
            self.interesting_fixpoints.append(net.get_weights())


class MixedFixpointExperiment(FixpointExperiment):

    def run_net(self, net, trains_per_application=100, step_limit=100, run_id=0):
        """Alternate one self_attack with `trains_per_application` train epochs (code/experiment.py:96-109)."""
        i = 0
        while i < step_limit and not net.is_diverged() and not net.is_fixpoint():
            net.self_attack()
            with tqdm(postfix=["Loss", dict(value=0)]) as bar:
                for _ in range(trains_per_application):
                    loss = net.compiled().train()
                    bar.postfix[1]["value"] = loss
                    bar.update()
            i += 1
            if run_id:
                net.save_state()
        self.count(net)


class SoupExperiment(Experiment):
    pass


class IdentLearningExperiment(Experiment):

    def __init__(self):
        super(IdentLearningExperiment, self).__init__(name=self.__class__.__name__)
