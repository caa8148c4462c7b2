"""The EEM driver loop of the reference's example programs as a function.

The reference has no driver class: every `examples/*/main.py` spells the same loop
(examples/bars-test/main.py:158-170, feature-learning/main.py, image-denoising/main.py:150-177):

    for e in range(no_epochs):
        F, S_nunique, S_sub, theta = model.step(theta, my_suff_stat, my_data[, do_reconstruction])
        dlog.append_all({"F": F, "S_nunique": S_nunique, "S_sub": S_sub, **theta})
        visualizer.process_epoch(epoch=e + 1, F=F, theta=theta)

`train` is that loop with the same call order and the same values handed to the log and the per-epoch
hook, plus what the reference lacks: a resumable checkpoint every `checkpoint_every`
epochs.  Everything heavy happens inside `model.step` (CUDA); this file is host orchestration only.
"""
import time

from . import checkpoint as _ck

# keys `model.step` adds to theta for its own use (E_step_precompute, bsc.py:100-125 / sssc.py:329-366);
# the reference logs them too, `log_internal=False` keeps the log to the learned parameters
_INTERNAL = ("pre1", "pil_bar", "piH", "ljc", "sigma2_inv")


def train(model, model_params, my_suff_stat, my_data, no_epochs, dlog=None, do_reconstruction=False,
          on_epoch=None, checkpoint_path=None, checkpoint_every=0, start_epoch=0, log_internal=True):
    """Run epochs `start_epoch .. no_epochs-1`; returns `(theta, history)` with
    `history = {"F": [...], "S_nunique": [...], "S_sub": [...], "seconds": [...]}`.

    dlog        anything with DataLog's `append_all(dict)` (evo/utils/datalog.py; `checkpoint.NpzLog` here)
    on_epoch    callable(epoch=<1-based>, F=, theta=) -- the visualizer hook of the examples
    checkpoint_path / checkpoint_every   write `save_checkpoint(path, theta, my_suff_stat, {"epoch": e + 1})`
                after every `checkpoint_every`-th epoch and after the last one (rank-local file: pass a
                per-rank path); resume with `load_checkpoint` and `start_epoch=extra["epoch"]`.
    """
    theta = model_params
    hist = {"F": [], "S_nunique": [], "S_sub": [], "seconds": []}
    for e in range(int(start_epoch), int(no_epochs)):
        t0 = time.time()
        F, S_nunique, S_sub, theta = model.step(theta, my_suff_stat, my_data, do_reconstruction=do_reconstruction)
        hist["F"].append(F)
        hist["S_nunique"].append(S_nunique)
        hist["S_sub"].append(S_sub)
        hist["seconds"].append(time.time() - t0)
        if dlog is not None:
            row = {"F": F, "S_nunique": S_nunique, "S_sub": S_sub}
            row.update({k: v for k, v in theta.items() if log_internal or k not in _INTERNAL})
            dlog.append_all(row)
        if on_epoch is not None:
            on_epoch(epoch=e + 1, F=F, theta=theta)
        last = e + 1 == int(no_epochs)
        if checkpoint_path and ((checkpoint_every and (e + 1) % int(checkpoint_every) == 0) or last):
            _ck.save_checkpoint(checkpoint_path, theta, my_suff_stat, extra={"epoch": e + 1})
    return theta, hist
