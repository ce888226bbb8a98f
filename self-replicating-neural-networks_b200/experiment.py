"""experiment -- host-side stand-in for the reference's code/experiment.py.

What lives here is bookkeeping, and it stays plain Python as in the reference: an experiment owns a directory
`experiments/exp-<name>-<id>-<iteration>/` (relative to the working directory), a list of log messages and
dill dumps of whatever the script hands to save().  The directory layout, the file names (`log.txt`,
`<key>.dill`, `experiment.dill`) and the attributes that end up inside the pickles (experiment_id,
experiment_name, next_iteration, log_messages, historical_particles, counters, ...) are kept, because the
reference's plotting scripts read them back.

The two hot loops of this module are NOT Python here: FixpointExperiment.run_net (code/experiment.py:70-77) and
count (:79-91) are one kernel launch each -- the stop rule, the per-step trajectory and the 5-way classification
run on the device (srnn_self_apply with SRNN_RUN_UNTIL, srnn_classify).

Names re-exported for `from experiment import *`, as the setups expect: os, time, dill, tqdm, copy.
"""
import copy
import os
import time

import dill
from tqdm import tqdm

__all__ = ["os", "time", "dill", "tqdm", "copy", "Experiment", "FixpointExperiment", "MixedFixpointExperiment",
           "SoupExperiment", "IdentLearningExperiment"]

_CLASS_LABELS = ("divergent", "fix_zero", "fix_other", "fix_sec", "other")
_UNPICKLED_ATTRS = ("particles", "historical_particles")     # live nets do not pickle; their `states` lists do


def _dill_to(path, obj):
    with open(path, "wb") as handle:
        dill.dump(obj, handle)


class Experiment:
    """`with Experiment('name') as exp:` creates the directory on entry and writes `experiment.dill` + `log.txt`
    on exit (also when the block raises, like the reference)."""

    def __init__(self, name=None, ident=None):
        self.experiment_id = "%s_%s" % (ident or "", time.time())
        self.experiment_name = name if name else "unnamed_experiment"
        self.next_iteration = 0
        self.log_messages = []
        self.historical_particles = {}

    # -- loading ---------------------------------------------------------------------------------------
    @staticmethod
    def from_dill(path):
        with open(path, "rb") as handle:
            return dill.load(handle)

    # -- context manager -------------------------------------------------------------------------------
    def _directory_name(self):
        return "exp-%s-%s-%s" % (self.experiment_name, self.experiment_id, self.next_iteration)

    def __enter__(self):
        self.dir = os.path.join("experiments", self._directory_name())
        os.makedirs(self.dir)            # an existing directory is an error, as in the reference
        print("** created %s **" % self.dir)
        return self

    def __exit__(self, exc_type, exc_value, traceback):
        self.save(experiment=self.without_particles())
        self.save_log()
        self.next_iteration += 1

    # -- logging ---------------------------------------------------------------------------------------
    def log(self, message, **print_kwargs):
        self.log_messages.append(message)
        print(message, **print_kwargs)

    def save_log(self, log_name="log"):
        target = os.path.join(self.dir, log_name + ".txt")
        with open(target, "w") as handle:
            handle.writelines(str(message) + "\n" for message in self.log_messages)

    # -- pickling --------------------------------------------------------------------------------------
    def __copy__(self):
        clone = Experiment(name=self.experiment_name)
        clone.__dict__ = {key: value for key, value in vars(self).items() if key not in _UNPICKLED_ATTRS}
        return clone

    def without_particles(self):
        """A shallow copy whose historical_particles maps uid -> list of state dicts."""
        clone = copy.copy(self)
        clone.historical_particles = {uid: particle.states for uid, particle in self.historical_particles.items()}
        return clone

    def save(self, **named_objects):
        for key, obj in named_objects.items():
            _dill_to(os.path.join(self.dir, key + ".dill"), obj)


class FixpointExperiment(Experiment):
    """Adds the class counters and the two loops every fixpoint script uses."""

    def __init__(self, **kwargs):
        kwargs["name"] = kwargs.get("name", type(self).__name__)
        super().__init__(**kwargs)
        self.counters = {label: 0 for label in _CLASS_LABELS}
        self.interesting_fixpoints = []

    def run_net(self, net, step_limit=100, run_id=0):
        """Self-apply `net` until it diverges, becomes a fixpoint or `step_limit` is reached; log a state per step
        when `run_id` is truthy; then classify it.  Engine nets do all of that in one fused launch."""
        fused = getattr(net, "_engine_run_until", None)
        if fused is not None and net._fused():
            fused(step_limit, log_states=bool(run_id))
        else:                                   # user subclass with its own operator: the generic loop
            steps_done = 0
            while steps_done < step_limit and not net.is_diverged() and not net.is_fixpoint():
                net.self_attack()
                steps_done += 1
                if run_id:
                    net.save_state(time=steps_done)
        self.count(net)

    def _label_of(self, net):
        fused = getattr(net, "_engine_class_name", None)
        if fused is not None and net._fused():
            return fused()
        if net.is_diverged():
            return "divergent"
        if net.is_fixpoint():
            return "fix_zero" if net.is_zero() else "fix_other"
        return "fix_sec" if net.is_fixpoint(2) else "other"

    def count(self, net):
        label = self._label_of(net)
        self.counters[label] += 1
        if label == "fix_other":
            self.interesting_fixpoints.append(net.get_weights())


class MixedFixpointExperiment(FixpointExperiment):

    def run_net(self, net, trains_per_application=100, step_limit=100, run_id=0):
        """Like run_net, but every self-application is followed by `trains_per_application` train() epochs
        (code/experiment.py:96-109); the tqdm bar shows the running loss as the reference's does."""
        rounds = 0
        while rounds < step_limit and not net.is_diverged() and not net.is_fixpoint():
            net.self_attack()
            with tqdm(postfix=["Loss", dict(value=0)]) as bar:
                for _ in range(trains_per_application):
                    bar.postfix[1]["value"] = net.compiled().train()
                    bar.update()
            rounds += 1
            if run_id:
                net.save_state()
        self.count(net)


class SoupExperiment(Experiment):
    """Plain Experiment; the reference uses the subclass name for the directory only."""


class IdentLearningExperiment(Experiment):

    def __init__(self):
        super().__init__(name=type(self).__name__)
