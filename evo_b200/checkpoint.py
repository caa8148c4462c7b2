"""Resumable dump of the training state and a per-key append log.

The reference cannot resume: its `DataLog` (evo/utils/datalog.py) streams values into a
PyTables HDF5 file through `AutoTable` (one extendable array per key, evo/utils/autotable.py:232-270) and
the K sets live only in process memory.  PyTables is not available here, so both go to `.npz`:

  save_checkpoint / load_checkpoint   theta + the bit-packed K sets [N][S][Hw] + lpj [N][S+S_perm] + every
                                      container field the E-step reads, including the Philox seed, the
                                      shard's global offset and the iteration counter -- a resumed run
                                      draws the same random streams and reproduces the uninterrupted
                                      trajectory bit for bit (tests/test_checkpoint_gpu.py);
  NpzLog                              `append(key, value)` / `append_all(dict)` / `assign(key, value)` with
                                      DataLog's call shapes (examples/*/main.py), flushed as one array
                                      per key stacked along axis 0 -- the layout AutoTable gives an HDF5
                                      reader (`f[key][i]` = i-th appended value).
"""
import os

import numpy as np
import torch

from . import variational as V

_FORMAT = 1
_SCALARS = ("H", "S", "S_perm", "n_parents", "n_children", "n_generations", "Mprime", "seed", "n_offset", "iteration")


def save_checkpoint(path, model_params, my_suff_stat, extra=None):
    """One `.npz` per rank (the K sets are this rank's shard).  `extra`: dict of small arrays / scalars."""
    ss = my_suff_stat
    out = {"format": np.int64(_FORMAT)}
    for k, v in model_params.items():
        out["theta/" + k] = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)
    out["kstates/ss_packed"] = dict.__getitem__(ss, "ss_dev").cpu().numpy()
    out["kstates/lpj"] = dict.__getitem__(ss, "lpj_dev").cpu().numpy()
    for k in _SCALARS:
        out["kstates/" + k] = np.int64(ss[k])
    out["kstates/bitflip_prob"] = np.float64(np.nan if ss["bitflip_prob"] is None else ss["bitflip_prob"])
    out["kstates/parent_selection"] = np.str_(ss["parent_selection"].name)
    out["kstates/mutation_algorithm"] = np.str_(ss["mutation_algorithm"].name)
    for k in ("background", "allzero", "singletons"):
        out["kstates/permanent/" + k] = np.int64(int(bool(ss["permanent"][k])))
    out["kstates/incl"] = np.asarray(ss["incl"], dtype=bool)
    for k, v in (extra or {}).items():
        out["extra/" + k] = np.asarray(v)
    tmp = path + ".tmp.npz"
    np.savez(tmp, **out)
    os.replace(tmp, path)  # a crash while writing leaves the previous checkpoint intact


def load_checkpoint(path, device=None):
    """-> (model_params, my_suff_stat, extra).  Parameters come back as NumPy arrays / Python floats (the
    model classes accept both); the container is ready for `model.step`."""
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    device = torch.device(device)
    with np.load(path, allow_pickle=False) as f:
        if int(f["format"]) != _FORMAT:
            raise ValueError("unknown checkpoint format %d" % int(f["format"]))
        theta, extra = {}, {}
        for k in f.files:
            if k.startswith("theta/"):
                v = f[k]
                theta[k[6:]] = float(v) if v.ndim == 0 else v
            elif k.startswith("extra/"):
                extra[k[6:]] = f[k]
        c = V.KStates()
        bf = float(f["kstates/bitflip_prob"])
        c.update({
            "ss_dev": torch.as_tensor(f["kstates/ss_packed"], device=device),
            "lpj_dev": torch.as_tensor(f["kstates/lpj"], device=device),
            "permanent": {k: bool(int(f["kstates/permanent/" + k])) for k in ("background", "allzero", "singletons")},
            "incl": f["kstates/incl"],
            "parent_selection": V._SELECTION[str(f["kstates/parent_selection"])],
            "mutation_algorithm": V._MUTATION[str(f["kstates/mutation_algorithm"])],
            "bitflip_prob": None if np.isnan(bf) else bf,
            "inject": None,
        })
        for k in _SCALARS:
            c[k] = int(f["kstates/" + k])
        Hm = c["H"] - 1 if c["permanent"]["background"] else c["H"]
        c["sm"] = V.all_states(Hm) if Hm < 12 else None
    return theta, c, extra


class NpzLog:
    """Append log with DataLog's call shapes; `flush()` rewrites `<path>` with one stacked array per key."""

    def __init__(self, path):
        self.path = path
        self._rows = {}
        self._fixed = {}

    def append(self, key, value):
        self._rows.setdefault(key, []).append(np.asarray(value.detach().cpu().numpy() if isinstance(value, torch.Tensor) else value))

    def append_all(self, values):
        for k, v in values.items():
            self.append(k, v)

    def assign(self, key, value):
        self._fixed[key] = np.asarray(value.detach().cpu().numpy() if isinstance(value, torch.Tensor) else value)

    def flush(self):
        out = dict(self._fixed)
        for k, rows in self._rows.items():
            out[k] = np.stack(rows, axis=0)
        tmp = self.path + ".tmp.npz"
        np.savez(tmp, **out)
        os.replace(tmp, self.path)

    close = flush
