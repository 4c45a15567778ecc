"""The training driver: hooks + monitored loop (reference main.py:111-156).

``MonitoredTrainingSession`` gives the reference: restore-or-init, periodic checkpoints by the chief, hook dispatch,
``should_stop()``.  :class:`MonitoredTrainingLoop` provides the same contract around an engine (native or torch):

* :class:`StopAtStepHook` (``last_step`` is a *global* step budget, main.py:137),
* :class:`NanLossHook` (``NanTensorHook(loss)``, main.py:138) — reads the device-side NaN flag, every
  ``nan_check_every`` steps and at the end, so that the hot loop is not serialised by a host read per step,
* :class:`LoggerHook` (the chief's ``_LoggerHook``, main.py:111-135): when ``global_step`` passes the next multiple of
  ``test_interval`` it writes the merged train summaries and then evaluates a *fresh* batch for the "test" summaries
  (the reference's extra ``session.run(test_summaries)``).
"""
from __future__ import annotations

import math
import time

import torch

from .monitor import DLMAO_TEST_SUMMARIES, DLMAO_TRAIN_SUMMARIES


class NanLossError(RuntimeError):
    """Raised when the loss becomes NaN/Inf (tf.train.NanLossDuringTrainingError)."""


class RunContext:
    def __init__(self, loop: "MonitoredTrainingLoop"):
        self.loop = loop
        self.stop_requested = False

    def request_stop(self) -> None:
        self.stop_requested = True


class Hook:
    def begin(self, loop) -> None: ...
    def before_run(self, ctx: RunContext) -> None: ...
    def after_run(self, ctx: RunContext) -> None: ...
    def end(self, loop) -> None: ...


class StopAtStepHook(Hook):
    def __init__(self, last_step: int):
        self.last_step = int(last_step)

    def begin(self, loop) -> None:
        if loop.global_step >= self.last_step:
            loop.ctx.request_stop()

    def after_run(self, ctx: RunContext) -> None:
        if ctx.loop.global_step >= self.last_step:
            ctx.request_stop()


class NanLossHook(Hook):
    """``check_every``: steps between polls.  Both engines keep a STICKY NaN flag (the native one on the device, set by the
    loss kernel of every step), so a sparse poll only delays the stop, it cannot miss the event.  Every poll of the native
    engine is a device->host read that drains the queue of launched steps, and with several ranks also a collective of
    the control plane: every 100 steps (15 ms) that costs well under 1 %; every 10 steps it was 5-10 % of a 150 us step."""

    def __init__(self, check_every: int = 100, fail_on_nan_loss: bool = True):
        self.check_every, self.fail = max(1, check_every), fail_on_nan_loss

    def _check(self, loop) -> None:
        m = loop.engine.metrics()
        loss = m.get("total_loss", m.get("cross_entropy", 0.0))
        bad = bool(m.get("nan")) or (isinstance(loss, float) and (math.isnan(loss) or math.isinf(loss)))
        # synchronous replicas take this decision together: a rank that stopped alone would leave its peers waiting in the
        # next step's cross-GPU barrier until the watchdog kills them (every rank polls on the same local step)
        if loop.any_rank(bad):
            if self.fail:
                raise NanLossError(f"NaN loss during training at global step {m.get('global_step')}")
            loop.ctx.request_stop()

    def after_run(self, ctx: RunContext) -> None:
        if ctx.loop.local_steps % self.check_every == 0:
            self._check(ctx.loop)

    def end(self, loop) -> None:
        self._check(loop)


class FaultInjectionHook(Hook):
    """Test hook for the failure path: ``DTM_FAULT_INJECT="<rank>:<local step>[:<attempt>]"`` makes that worker
    die abruptly (``os._exit(43)``: no clean-up, no goodbye to its peers - what a crashed process looks like) after that
    many local steps, on that restart attempt (default 0, the first run) of a supervised job."""

    EXIT_CODE = 43

    def __init__(self, spec: str, rank: int, attempt: int, out=print):
        parts = [int(x) for x in spec.split(":")]
        self.rank, self.step = parts[0], parts[1]
        self.attempt = parts[2] if len(parts) > 2 else 0
        self.active = rank == self.rank and attempt == self.attempt
        self.out = out

    def after_run(self, ctx: RunContext) -> None:
        if self.active and ctx.loop.local_steps >= self.step:
            import os
            import sys

            self.out(f"[fault-injection] worker {self.rank} dies after local step {ctx.loop.local_steps}")
            sys.stdout.flush()
            os._exit(self.EXIT_CODE)


class StepCounterHook(Hook):
    """``global_step/sec`` every ``every_n_steps`` global steps into the event file of ``train_dir`` itself — the default
    ``StepCounterHook`` that ``MonitoredTrainingSession`` installs on the chief.  ``images/sec`` is
    written next to it when the global batch size is known (the reference never reports throughput: main.py:149-154).
    Host wall clock over asynchronous steps, i.e. the steady-state rate, like TensorFlow's hook."""

    def __init__(self, writer, every_n_steps: int = 100, global_batch: int | None = None):
        self.writer, self.every, self.global_batch = writer, max(1, every_n_steps), global_batch
        self._t = self._step = None
        self.last_rate = None

    def begin(self, loop) -> None:
        self._t, self._step = time.time(), loop.global_step

    def after_run(self, ctx: RunContext) -> None:
        from ..formats.events import scalar_value

        gs = ctx.loop.global_step
        if gs - self._step < self.every:
            return
        now = time.time()
        rate = (gs - self._step) / max(now - self._t, 1e-9)
        self.last_rate = rate
        self.writer.add_summary(scalar_value("global_step/sec", rate), gs)
        if self.global_batch:
            self.writer.add_summary(scalar_value("images/sec", rate * self.global_batch), gs)
        self._t, self._step = now, gs

    def end(self, loop) -> None:
        self.writer.flush()


class LoggerHook(Hook):
    def __init__(self, monitor, writer, test_interval: int):
        self.monitor, self.writer, self.test_interval = monitor, writer, max(1, int(test_interval))
        self._next = self.test_interval
        self._trigger = False

    def begin(self, loop) -> None:
        self._next = (loop.global_step // self.test_interval + 1) * self.test_interval
        self._trigger = False

    def after_run(self, ctx: RunContext) -> None:
        loop = ctx.loop
        gs = loop.global_step
        if self._trigger:
            # the step that just ran is the one whose summaries the reference fetches (main.py:120-122, :131-135)
            self._trigger = False
            self.write(loop, gs)
        if gs >= self._next:
            self._trigger = True
            self._next += self.test_interval

    def write(self, loop, gs: int) -> None:
        m = loop.engine.metrics()
        tensors = loop.summary_tensors()
        train = self.monitor.collect(DLMAO_TRAIN_SUMMARIES, tensors, {"loss": m["total_loss"], "accuracy": m["accuracy"]})
        self.writer.add_summary(train, gs)
        loss, acc = loop.evaluate_fresh_batch()
        test = self.monitor.collect(DLMAO_TEST_SUMMARIES, tensors, {"loss": loss, "accuracy": acc})
        self.writer.add_summary(test, gs)

    def end(self, loop) -> None:
        self.writer.flush()


class MonitoredTrainingLoop:
    """``with MonitoredTrainingLoop(...) as sess: while not sess.should_stop(): sess.run_step()``."""

    def __init__(self, engine, source, *, hooks=(), is_chief=True, saver=None, task_id=0, feeder=None, log_every=100,
                 verbose_steps=False, out=print, group=None, save_poll_every=100):
        self.engine, self.source = engine, source
        self.hooks = list(hooks)
        self.is_chief, self.saver, self.task_id = is_chief, saver, task_id
        self.feeder = feeder
        self.ctx = RunContext(self)
        self.local_steps = 0
        self.start_step = int(engine.metrics()["global_step"])
        self.log_every, self.verbose_steps, self.out = log_every, verbose_steps, out
        self._t0 = None
        # A sharded engine (fused multi-GPU update) needs every rank to take part in assembling a checkpoint, so the
        # chief's time-based decision is broadcast at fixed step intervals and all ranks act on it together.
        self.group = group        # torch.distributed process group or a control plane (parallel/control_plane.py)
        self._cp = None
        self.collective_saves = bool(getattr(engine, "sharded", False)) and group is not None and saver is not None
        self.save_poll_every = max(1, save_poll_every)

    # the global step advances by exactly one per synchronous step, so the host can track it without a device read
    @property
    def global_step(self) -> int:
        return self.start_step + self.local_steps

    def __enter__(self):
        for h in self.hooks:
            h.begin(self)
        self._t0 = time.time()
        return self

    def should_stop(self) -> bool:
        return self.ctx.stop_requested

    def run_step(self) -> None:
        for h in self.hooks:
            h.before_run(self.ctx)
        if self.feeder is not None:
            self.feeder.step()
        else:
            images, labels = self.source.next()
            self.engine.stage_batch(0, images, labels)
            self.engine.train_step(0)
        self.local_steps += 1
        for h in self.hooks:
            h.after_run(self.ctx)
        if self.collective_saves:
            if self.local_steps % self.save_poll_every == 0 and self._agree(self.is_chief and self.saver.due()):
                self.saver.save(self.engine, write=self.is_chief)
        elif self.is_chief and self.saver is not None:
            self.saver.maybe_save(self.engine)
        if self.verbose_steps:           # the reference prints "training" and the step every iteration (main.py:149-150)
            self.out("training")
            self.out(self.local_steps - 1)
        if self.log_every and self.local_steps % self.log_every == 0:
            self.out("%d steps executed on worker %d." % (self.local_steps, self.task_id))

    def _plane(self):
        if self._cp is None and self.group is not None:
            from ..parallel.control_plane import as_plane

            self._cp = as_plane(self.group)
        return self._cp

    def any_rank(self, flag: bool) -> bool:
        """True on every rank as soon as ``flag`` is true on one of them (collective; a no-op without a group)."""
        cp = self._plane()
        if cp is None or cp.world_size <= 1:
            return bool(flag)
        return cp.any_true(flag)

    def _agree(self, flag: bool) -> bool:
        """The chief's decision, known to every rank (callers pass False on every other rank: one cheap flag reduction
        instead of a pickled broadcast)."""
        return self._plane().any_true(bool(flag))

    def summary_tensors(self) -> dict:
        eng = self.engine
        if hasattr(eng, "summary_tensors"):
            return eng.summary_tensors()
        return {"params": eng.params(), "activations": eng.activations() if hasattr(eng, "activations") else {},
                "gradients": eng.gradients()}

    def evaluate_fresh_batch(self) -> tuple[float, float]:
        images, labels = self.source.next()
        slot = 0 if self.feeder is None else 1 - (self.feeder._step % 2)
        if torch.cuda.is_available() and getattr(self.engine, "device", torch.device("cpu")).type == "cuda":
            torch.cuda.synchronize()
        self.engine.stage_batch(slot, images, labels)
        res = self.engine.evaluate(slot)
        if self.feeder is not None:
            self.feeder.invalidate(slot)
        return res

    def __exit__(self, exc_type, exc, tb):
        if exc_type is None:
            for h in self.hooks:
                h.end(self)
            if self.collective_saves:
                self.saver.save(self.engine, write=self.is_chief)
            elif self.is_chief and self.saver is not None:
                self.saver.maybe_save(self.engine, force=True)
        self.out("%d steps executed on worker %d." % (self.local_steps, self.task_id))
        return False
