"""Wall-clock tic/toc (reference: src/common/timer.py:3-19)."""
from datetime import datetime


class Timer:
    def __init__(self):
        self._tic = None

    def tic(self):
        self._tic = datetime.now()

    def toc(self):
        """Seconds since the last tic()."""
        return (datetime.now() - self._tic).total_seconds()
