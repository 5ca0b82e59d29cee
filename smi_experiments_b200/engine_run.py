"""Epoch / minibatch run driver (reference src/engine_run.py:11-70), the direct caller of
``engine.update`` in experiments/img_classifier.py.

``setup_run(engine, dataset) -> run(rng_key, num_epochs, batch_size, init_state) -> Result``.
Differences from the reference, all on the host side of the boundary: the per-step
``device_get(loss)`` (src/engine_run.py:56) is replaced by the engine's device-side loss ring --
one D2H copy of all losses when the run ends; the ASVGD temperature follows the cyclical schedule
per step (the reference's annealed ``body_fn`` is shadowed by the plain one -- SURVEY.md Appendix B,
bug B3 -- pass ``compat_b3=True`` to reproduce that).
"""
from collections import namedtuple

import torch

from .data import train_batches
from .steinvi import ASVGD

Result = namedtuple("Result", ["params", "state", "losses"])


def setup_run(engine, dataset, compat_b3=False, seed=None):
    def run(rng_key, num_epochs, batch_size, init_state=None):
        if num_epochs < 1:
            raise ValueError("num_epochs must be a positive integer.")
        n = len(dataset[0])
        steps_per_epoch = n // batch_size
        total = num_epochs * steps_per_epoch
        anneal = isinstance(engine, ASVGD) and not compat_b3
        if anneal:
            engine._begin_schedule(num_epochs)     # src/engine_run.py:17-19: the schedule length is num_epochs
        state, losses0 = init_state, []
        ring = torch.zeros(max(total, 1), dtype=torch.float32, device=engine.device)
        engine._ring = ring
        start = None
        try:
            for epoch in range(num_epochs):
                batches, n_total = train_batches(batch_size, dataset, None if seed is None else seed + epoch)
                for x, y in batches:
                    if state is None:
                        state = engine.init(rng_key, x, y, n=n_total, batch_size=batch_size)
                        losses0.append(engine.evaluate(state, x, y, n=n_total, batch_size=batch_size).reshape(1))
                        start = int(state.optim_state.step.item())
                        continue
                    if start is None:
                        start = int(engine._own(state).optim_state.step.item())
                    state, _ = engine.update(state, x, y, n=n_total, batch_size=batch_size)
            done = int(state.optim_state.step.item()) - start
            k = start % ring.numel()
            rolled = torch.roll(ring, -k) if k else ring
            losses = torch.cat(losses0 + [rolled[:done]])
        finally:
            engine._ring = None
            engine._graph = None
            if anneal:
                engine._end_schedule()
        return Result(engine.get_params(state), state, losses)

    return run
