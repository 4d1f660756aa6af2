"""The slice of the reference's key-value logger (``basic_utils/logger.py``)
that the update path feeds: ``logkv`` / ``logkv_mean`` / ``dumpkvs`` / ``log``
/ ``configure`` / ``get_current().dir``.  File formats, wandb and MPI averaging
are outside this repository's scope; when the reference's own logger module is
importable ``TrainLoop`` can be handed that instead (``TrainLoop(logger=...)``).
"""
import csv
import os
import sys
import tempfile
from collections import defaultdict


class Logger:
    CURRENT = None

    def __init__(self, dir=None, stdout=False):
        self.dir = dir
        self.name2val = defaultdict(float)
        self.name2cnt = defaultdict(int)
        self.stdout = stdout
        self._csv_path = os.path.join(dir, "progress.csv") if dir else None
        self._csv_keys = None

    def logkv(self, key, val):
        self.name2val[key] = val

    def logkv_mean(self, key, val):
        # running mean, same update rule as reference logger.py:353-356
        old, cnt = self.name2val[key], self.name2cnt[key]
        self.name2val[key] = old * cnt / (cnt + 1) + val / (cnt + 1)
        self.name2cnt[key] = cnt + 1

    def logkv_mean_n(self, key, mean, n):
        """``n`` calls of logkv_mean with samples whose mean is ``mean``, in one update."""
        if n <= 0:
            return
        old, cnt = self.name2val[key], self.name2cnt[key]
        self.name2val[key] = old * cnt / (cnt + n) + mean * n / (cnt + n)
        self.name2cnt[key] = cnt + n

    def dumpkvs(self):
        out = dict(self.name2val)
        if int(os.environ.get("LOCAL_RANK", "0")) == 0:
            if self.stdout:
                for k in sorted(out):
                    print(f"| {k:<24}| {out[k]!s:<16}|", file=sys.stdout)
            if self._csv_path:
                keys = sorted(out)
                new = self._csv_keys != keys
                self._csv_keys = keys
                with open(self._csv_path, "a", newline="") as f:
                    w = csv.writer(f)
                    if new:
                        w.writerow(keys)
                    w.writerow([out[k] for k in keys])
        self.name2val.clear()
        self.name2cnt.clear()
        return out

    def log(self, *args):
        if int(os.environ.get("LOCAL_RANK", "0")) == 0:
            print(*args)

    def get_dir(self):
        return self.dir


def configure(dir=None, stdout=False, **_ignored):
    if dir is None:
        dir = os.environ.get("OPENAI_LOGDIR") or tempfile.mkdtemp(prefix="dpk-log-")
    os.makedirs(dir, exist_ok=True)
    Logger.CURRENT = Logger(dir, stdout=stdout)
    return Logger.CURRENT


def get_current():
    if Logger.CURRENT is None:
        configure()
    return Logger.CURRENT


def logkv(key, val):
    get_current().logkv(key, val)


def logkv_mean(key, val):
    get_current().logkv_mean(key, val)


def logkv_mean_n(key, mean, n):
    get_current().logkv_mean_n(key, mean, n)


def dumpkvs():
    return get_current().dumpkvs()


def getkvs():
    return get_current().name2val


def log(*args):
    get_current().log(*args)


def get_dir():
    return get_current().get_dir()
