"""
StepSequence (pyrado/sampling/step_sequence.py:223-1047) -- the container `rollout()` and the samplers return -- and
the return helpers (:1049-1116).  Fields are stacked arrays with the step count as first dimension; `observations`,
`states` (and any field added with one extra element) hold `length + 1` rows.  The engine fills these arrays from the
time-major record buffer the rollout kernel writes (`from_records`), as numpy on the host or as torch views that stay
on the device.
"""
from typing import Iterable, Optional, Sequence, Union

import numpy as np
import torch

import simurlacra_b200 as pyrado


def _to_format(x, data_format: str):
    if data_format == "torch":
        if isinstance(x, torch.Tensor):
            return x
        return torch.as_tensor(np.asarray(x), dtype=torch.get_default_dtype())
    if isinstance(x, torch.Tensor):
        return x.detach().cpu().numpy()
    return np.asarray(x)


def _stack(seq, data_format: str):
    if isinstance(seq, (np.ndarray, torch.Tensor)):
        return _to_format(seq, data_format)
    if len(seq) > 0 and isinstance(seq[0], torch.Tensor):
        return _to_format(torch.stack(list(seq)), data_format)
    return _to_format(np.stack([np.asarray(s) for s in seq]), data_format)


class Step:
    """One step of a rollout: attribute access maps singular names to the row of the plural field, `next_<name>` to
    the following row (e.g. `step.observation`, `step.next_observation`, `step.reward`, `step.done`)."""

    def __init__(self, rollout: "StepSequence", index: int):
        object.__setattr__(self, "_rollout", rollout)
        object.__setattr__(self, "_index", index)

    def _resolve(self, key: str):
        ro, idx = self._rollout, self._index
        if key.startswith("next_"):
            key, idx = key[5:], idx + 1
        for cand in (key, key + "s"):
            if cand in ro.data_names:
                return cand, idx
        raise AttributeError(f"A step has no entry {key!r}")

    def __getattr__(self, key):
        if key == "done":
            return bool(self._rollout.done[self._index])
        name, idx = self._resolve(key)
        value = getattr(self._rollout, name)
        if isinstance(value, dict):
            return {k: v[idx] for k, v in value.items()}
        return value[idx]

    def __setattr__(self, key, value):
        name, idx = self._resolve(key)
        getattr(self._rollout, name)[idx] = value

    __getitem__ = __getattr__


class StepSequence(Sequence):
    """A sequence of steps; `rewards`, `observations` and `actions` are required, everything else is optional."""

    def __init__(self, *, complete: bool = True, rollout_info=None, data_format: Optional[str] = None,
                 done: Optional[np.ndarray] = None, continuous: bool = True, rollout_bounds=None, rewards: Sequence,
                 observations: Sequence, actions: Sequence, **data):
        if len(rewards) == 0 or len(observations) == 0 or len(actions) == 0:
            raise pyrado.ShapeErr(msg="StepSequence must not be empty!")
        self.rollout_info = rollout_info
        self.continuous = continuous
        self._data_names = []
        if data_format is None:
            is_t = lambda v: isinstance(v, torch.Tensor) or (isinstance(v, list) and len(v) and isinstance(v[0], torch.Tensor))
            data_format = "torch" if any(is_t(v) for v in data.values()) else "numpy"
        self._data_format = data_format
        self._data_names.append("rewards")
        self.__dict__["rewards"] = _stack(rewards, data_format)
        self.add_data("observations", observations)
        self.add_data("actions", actions)
        for name, value in data.items():
            self.add_data(name, value)
        if done is None:
            done = np.zeros(self.length, dtype=np.bool_)
            if complete and continuous:
                done[-1] = True
        else:
            done = np.asarray(done, dtype=np.bool_)
            assert done.shape[0] == self.length
        self.done = done
        if continuous:
            if rollout_bounds is None:
                rollout_bounds = [0, *(np.flatnonzero(done) + 1)]
                if not done[-1]:
                    rollout_bounds.append(self.length)
            else:
                rb = list(rollout_bounds)
                if any(not rb[i] < rb[i + 1] for i in range(len(rb) - 1)):
                    raise pyrado.ValueErr(msg="The lower rollout bound must be smaller than the upper for every rollout!")
                if rb[0] != 0:
                    raise pyrado.ValueErr(given=rb[0], eq_constraint="0")
                if rb[-1] != self.length:
                    raise pyrado.ValueErr(given=rb[-1], eq_constraint=self.length)
            self._rollout_bounds = np.array(rollout_bounds)
        else:
            self._rollout_bounds = None

    def __repr__(self) -> str:
        return (f"StepSequence[\n rewards=\n {self.rewards}\n observations=\n {self.observations}\n actions=\n "
                f"{self.actions}\n ]")

    # ------------------------------------------------------------------ properties
    @property
    def length(self) -> int:
        """Number of steps (the final observation / state is not counted)."""
        return len(self.rewards)

    def __len__(self):
        return self.length

    @property
    def data_format(self) -> str:
        return self._data_format

    @property
    def data_names(self) -> Sequence[str]:
        return self._data_names

    @property
    def rollout_bounds(self) -> np.ndarray:
        assert self.continuous
        return self._rollout_bounds

    @property
    def rollout_count(self) -> int:
        if not self.continuous:
            raise pyrado.ValueErr(msg="Sub-rollouts are only supported on continuous data.")
        return len(self._rollout_bounds) - 1

    @property
    def rollout_lengths(self):
        if not self.continuous:
            raise pyrado.ValueErr(msg="Sub-rollouts are only supported on continuous data.")
        return self._rollout_bounds[1:] - self._rollout_bounds[:-1]

    # ------------------------------------------------------------------ data management
    def _validate(self, name, value):
        n = len(value)
        if n not in (self.length, self.length + 1):
            raise pyrado.ShapeErr(msg=f"The data entry {name} has {n} rows, expected {self.length} or {self.length + 1}!")

    def add_data(self, name: str, value=None, item_shape: tuple = None, with_after_last: bool = False):
        """Add a field; `value` is a sequence / array of per-step entries, a list of dicts, or None (zeros)."""
        if name in self._data_names:
            raise pyrado.KeyErr(msg=f"Trying to add a duplicate data field for {name}!")
        if value is None:
            rows = self.length + int(with_after_last)
            shape = (rows, *(item_shape or ()))
            value = torch.zeros(shape) if self._data_format == "torch" else np.zeros(shape)
        elif isinstance(value, (list, tuple)) and len(value) and isinstance(value[0], dict):
            keys = value[0].keys()
            value = {k: _stack([v[k] for v in value], self._data_format) for k in keys}
            for k, v in value.items():
                self._validate(f"{name}.{k}", v)
        elif isinstance(value, dict):
            value = {k: _stack(v, self._data_format) for k, v in value.items()}
        else:
            value = _stack(value, self._data_format)
            self._validate(name, value)
        self._data_names.append(name)
        self.__dict__[name] = value

    def get_data_values(self, name: str, truncate_last: bool = False):
        """The field's array; `truncate_last` drops the extra final row of `length + 1` fields."""
        assert name in self._data_names
        entry = self.__dict__[name]
        if truncate_last:
            cut = lambda v: v[:-1] if len(v) == self.length + 1 else v
            entry = {k: cut(v) for k, v in entry.items()} if isinstance(entry, dict) else cut(entry)
        return entry

    def _map(self, fcn):
        for name in self._data_names:
            v = self.__dict__[name]
            self.__dict__[name] = {k: fcn(x) for k, x in v.items()} if isinstance(v, dict) else fcn(v)

    def numpy(self, data_type=None):
        self.convert("numpy", data_type)

    def torch(self, data_type=None):
        self.convert("torch", data_type)

    def convert(self, data_format: str, data_type=None):
        """Convert all fields in place to 'numpy' or 'torch' (optionally casting)."""
        if data_format not in ("torch", "numpy"):
            raise pyrado.ValueErr(given=data_format, eq_constraint="'torch' or 'numpy'")
        if data_format == "torch":
            self._map(lambda x: torch.as_tensor(x).to(dtype=data_type or torch.get_default_dtype())
                      if not isinstance(x, torch.Tensor) else (x if data_type is None else x.to(data_type)))
        else:
            self._map(lambda x: _to_format(x, "numpy") if data_type is None else _to_format(x, "numpy").astype(data_type))
        self._data_format = data_format

    # ------------------------------------------------------------------ slicing
    def __getitem__(self, index):
        if isinstance(index, (slice, Iterable)) and not isinstance(index, (str, bytes)):
            return self._slice(index)
        n = self.length
        if not -n <= index < n:
            raise IndexError(f"Step index {index} out of range for a rollout of length {n}")
        return Step(self, index % n)

    def __iter__(self):
        return (Step(self, i) for i in range(self.length))

    def _slice(self, index):
        """Every field is indexed with `index` (rows of `length + 1` fields included), like the reference (:409-437)."""
        continuous = isinstance(index, slice) and index.step in (None, 1)
        if not isinstance(index, slice):
            index = np.asarray(list(index))
        cut = lambda v: {k: x[index] for k, x in v.items()} if isinstance(v, dict) else v[index]
        data = {name: cut(self.__dict__[name]) for name in self._data_names}
        bounds = None
        if continuous and self.continuous:
            start, stop, _ = index.indices(self.length)
            bounds = [0, *[int(x) - start for x in self._rollout_bounds if start < x < stop], stop - start]
        return StepSequence(rollout_info=self.rollout_info, data_format=self._data_format, done=self.done[index],
                            continuous=continuous and self.continuous, rollout_bounds=bounds, **data)

    def get_rollout(self, index):
        """Sub-rollout(s) of a concatenated sequence by rollout index or slice of indices."""
        if not self.continuous:
            raise pyrado.ValueErr(msg="Sub-rollouts are only supported on continuous data.")
        b = self._rollout_bounds
        if isinstance(index, slice):
            lo, hi, step = index.indices(self.rollout_count)
            if step != 1:
                raise pyrado.ValueErr(msg="Rollout slices must be contiguous")
            return self[int(b[lo]):int(b[hi])]
        index = index % self.rollout_count
        return self[int(b[index]):int(b[index + 1])]

    def iterate_rollouts(self):
        if not self.continuous:
            raise pyrado.ValueErr(msg="Sub-rollouts are only supported on continuous data.")
        for i in range(self.rollout_count):
            yield self.get_rollout(i)

    # ------------------------------------------------------------------ batching
    def sample_w_next(self, batch_size: int) -> tuple:
        """A random batch of steps together with their successors (:658-674)."""
        if not self.length >= 2:
            raise pyrado.ValueErr(given=self.length, ge_constraint="2")
        pool = np.flatnonzero(~self.done[:-1])  # a step that ended an episode has no successor
        pool = pool if len(pool) else np.arange(self.length - 1)
        idx = np.random.choice(pool, min(batch_size, len(pool)), replace=False)
        return self[idx], self[idx + 1]

    def split_ordered_batches(self, batch_size: int = None, num_batches: int = None):
        """Consecutive mini-batches of `batch_size` steps (or `num_batches` of equal size), (:676-704)."""
        if (batch_size is None) == (num_batches is None):
            raise pyrado.ValueErr(msg="Either batch_size or num_batches must not be None, but not both or none!")
        if num_batches is not None:
            batch_size = int(np.ceil(self.length / num_batches))
        if batch_size >= self.length:
            yield self
        else:
            for b in range(int(np.ceil(self.length / batch_size))):
                yield self[b * batch_size: min((b + 1) * batch_size, self.length)]

    def split_shuffled_batches(self, batch_size: int, complete_rollouts: bool = False):
        """Randomly composed mini-batches; with `complete_rollouts` whole sub-rollouts are kept together (:706-746)."""
        if batch_size >= self.length:
            yield self
        elif complete_rollouts and self.continuous:
            order = np.random.permutation(self.rollout_count)
            batch, n = [], 0
            for i in order:
                ro = self.get_rollout(int(i))
                batch.append(ro)
                n += ro.length
                if n >= batch_size:
                    yield StepSequence.concat(batch, truncate_last=True)
                    batch, n = [], 0
            if batch:
                yield StepSequence.concat(batch, truncate_last=True)
        else:
            perm = np.random.permutation(self.length)
            for b in range(int(np.ceil(self.length / batch_size))):
                yield self[perm[b * batch_size: (b + 1) * batch_size]]

    # ------------------------------------------------------------------ returns
    def undiscounted_return(self) -> float:
        if not len(self._rollout_bounds) == 2:
            raise pyrado.ShapeErr(msg="The StepSequence must be a single continuous rollout.")
        return self.rewards.sum()

    def discounted_return(self, gamma: float):
        if not len(self._rollout_bounds) == 2:
            raise pyrado.ShapeErr(msg="The StepSequence must be a single continuous rollout.")
        if not 0 <= gamma <= 1:
            raise pyrado.ValueErr(given=gamma, ge_constraint="0", le_constraint="1")
        if self.data_format == "torch":
            return torch.dot(self.rewards, (gamma ** torch.arange(self.length)).to(self.rewards))
        return np.dot(self.rewards, gamma ** np.arange(self.length))

    # ------------------------------------------------------------------ construction helpers
    @classmethod
    def concat(cls, parts: Sequence["StepSequence"], data_format: Optional[str] = None, truncate_last: bool = True):
        """Concatenate step sequences into one, dropping each part's final observation / state row (:776-825)."""
        data_format = data_format or parts[0].data_format

        def cat(vals):
            if isinstance(vals[0], dict):
                return {k: cat([v[k] for v in vals]) for k in vals[0]}
            vals = [_to_format(v, data_format) for v in vals]
            return torch.cat(vals) if data_format == "torch" else np.concatenate(vals)

        data = {name: cat([ro.get_data_values(name, truncate_last) for ro in parts]) for name in parts[0].data_names}
        continuous = all(ro.continuous for ro in parts)
        bounds = None
        if continuous:
            bounds, acc = [0], 0
            for ro in parts:
                bounds.extend(int(b) + acc for b in ro.rollout_bounds[1:])
                acc += int(ro.rollout_bounds[-1])
        infos = [p.rollout_info for p in parts]
        same = all(i is infos[0] for i in infos) or all(_info_equal(i, infos[0]) for i in infos)
        return StepSequence(data_format=data_format, done=np.concatenate([ro.done for ro in parts]),
                            continuous=continuous, rollout_bounds=bounds, rollout_info=infos[0] if same else None, **data)

    @classmethod
    def pad(cls, rollout: "StepSequence", len_to_pad_to: int, pad_value: Union[int, float] = 0):
        """Pad a rollout in place with `pad_value` rows up to `len_to_pad_to` steps (:978-1047)."""
        if len_to_pad_to <= rollout.length:
            return
        extra = len_to_pad_to - rollout.length

        def pad_one(v):
            if isinstance(v, dict):
                return {k: pad_one(x) for k, x in v.items()}
            shape = (extra, *v.shape[1:])
            if isinstance(v, torch.Tensor):
                return torch.cat([v, torch.full(shape, pad_value, dtype=v.dtype, device=v.device)])
            return np.concatenate([v, np.full(shape, pad_value, dtype=v.dtype)])

        for name in rollout._data_names:
            rollout.__dict__[name] = pad_one(rollout.__dict__[name])
        rollout.done = np.concatenate([rollout.done, np.full(extra, bool(pad_value), dtype=np.bool_)])
        if rollout.continuous:
            rollout._rollout_bounds = np.array([*rollout._rollout_bounds[:-1], len_to_pad_to])

    @classmethod
    def from_records(cls, fields: dict, length: int, dt: float, rollout_info=None, done_last: bool = True,
                     data_format: str = "numpy") -> "StepSequence":
        """One episode from the columns of the kernel's record buffer: `fields` maps field name -> array whose first
        `length` (or `length + 1`) rows belong to the episode (engine extension)."""
        done = np.zeros(length, dtype=np.bool_)
        done[-1] = done_last
        time = np.arange(length + 1) * dt
        return cls(rewards=fields["rewards"], observations=fields["observations"], actions=fields["actions"],
                   done=done, rollout_info=rollout_info, data_format=data_format, complete=done_last,
                   time=time if data_format == "numpy" else torch.as_tensor(time),
                   **{k: v for k, v in fields.items() if k not in ("rewards", "observations", "actions")})


def _from_views(cls, fields: dict, length: int, time, rollout_info, done_last: bool, data_format: str) -> "StepSequence":
    """One complete episode whose fields are already arrays / tensors of the right format and length (views into the
    kernel's record buffer): no stacking, no validation, no copies -- what `ParallelRolloutSampler.sample()` builds for
    every rollout index (engine extension; `StepSequence.from_records` is the checked variant)."""
    self = cls.__new__(cls)
    self.rollout_info = rollout_info
    self.continuous = True
    self._data_format = data_format
    self._data_names = list(fields.keys()) + ["time"]
    self.__dict__.update(fields)
    self.__dict__["time"] = time
    done = np.zeros(length, dtype=np.bool_)
    done[-1] = done_last
    self.done = done
    self._rollout_bounds = np.array([0, length])
    return self


StepSequence.from_views = classmethod(_from_views)


def _info_equal(a, b) -> bool:
    try:
        if isinstance(a, dict) and isinstance(b, dict):
            return a.keys() == b.keys() and all(_info_equal(a[k], b[k]) for k in a)
        return bool(np.all(a == b))
    except Exception:
        return False


def discounted_reverse_cumsum(data, gamma: float):
    """Reverse cumulative sum of discounted values, y[t] = x[t] + gamma * y[t+1] (:1049-1060)."""
    out = data.clone() if isinstance(data, torch.Tensor) else np.array(data, dtype=np.float64, copy=True)
    for t in range(len(out) - 2, -1, -1):
        out[t] = out[t] + gamma * out[t + 1]
    return out


def discounted_value(rollout: StepSequence, gamma: float):
    """Discounted return-to-go of every step of ONE rollout (:1063-1072)."""
    rewards = [step.reward for step in rollout]
    stacked = torch.stack([torch.as_tensor(r) for r in rewards]) if rollout.data_format == "torch" else np.stack(rewards)
    return discounted_reverse_cumsum(stacked, gamma)


def discounted_values(rollouts: Sequence[StepSequence], gamma: float, data_format: Optional[str] = "torch"):
    """Discounted returns-to-go of several rollouts with the reference's return shapes (:1075-1090): 'torch' gives ONE
    1-D tensor (the rollouts concatenated), 'numpy' gives `np.array([...])` with one row per rollout ([n_ro, len] for
    equally long rollouts)."""
    if data_format == "torch":
        return torch.cat([torch.as_tensor(discounted_value(ro, gamma)).to(torch.get_default_dtype()).reshape(-1)
                          for ro in rollouts])
    if data_format == "numpy":
        vals = [np.asarray(_to_format(discounted_value(ro, gamma), "numpy"), dtype=np.float64).reshape(-1) for ro in rollouts]
        if len({len(v) for v in vals}) > 1:  # ragged: numpy >= 1.24 needs the object dtype spelt out
            out = np.empty(len(vals), dtype=object)
            for i, v in enumerate(vals):
                out[i] = v
            return out
        return np.array(vals)
    raise pyrado.ValueErr(given=data_format, eq_constraint="'torch' or 'numpy'")


def gae_returns(rollout: StepSequence, gamma: float = 0.99, lamb: float = 0.95):
    """Generalized advantage estimates of one rollout that carries a `values` field (:1093-1116)."""
    if "values" not in rollout.data_names:
        raise pyrado.KeyErr(keys="values", container=rollout)
    v = np.asarray(_to_format(rollout.values, "numpy"), dtype=np.float64).reshape(-1)
    r = np.asarray(_to_format(rollout.rewards, "numpy"), dtype=np.float64).reshape(-1)
    n = rollout.length
    nxt = np.zeros(n)
    nxt[:-1] = v[1:n]
    if len(v) == n + 1:
        nxt[-1] = v[n]
    nxt[rollout.done] = 0.0  # no bootstrapping across the end of an episode
    deltas = r + gamma * nxt - v[:n]
    adv = np.array(deltas, copy=True)
    for t in range(n - 2, -1, -1):
        if not rollout.done[t]:
            adv[t] = deltas[t] + gamma * lamb * adv[t + 1]
    return torch.as_tensor(adv, dtype=torch.get_default_dtype()) if rollout.data_format == "torch" else adv
