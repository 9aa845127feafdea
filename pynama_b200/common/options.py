"""Tiny stand-in for the PETSc options database used by the reference (`petsc4py.init(sys.argv)`,
`PETSc.Options().getString(...)`, `setFromOptions()`: src/run_case.py:5-10,165-167, src/cases/base_problem.py:29,
src/solver/ksp_solver.py:17, src/solver/ts_solver.py:26).  Accepts the same `-key value` spellings."""
import sys


class Options(object):
    _db = None

    def __init__(self, argv=None):
        if argv is not None or Options._db is None:
            Options._db = self._parse(sys.argv[1:] if argv is None else list(argv))

    @staticmethod
    def _parse(args):
        db, i = {}, 0
        while i < len(args):
            a = args[i]
            if a.startswith("-") and not _is_number(a):
                key = a.lstrip("-")
                if i + 1 < len(args) and (not args[i + 1].startswith("-") or _is_number(args[i + 1])):
                    db[key] = args[i + 1]
                    i += 2
                    continue
                db[key] = True
            i += 1
        return db

    @classmethod
    def init(cls, argv):
        cls._db = cls._parse(list(argv)[1:])

    @classmethod
    def setValue(cls, key, value):
        if cls._db is None:
            cls._db = {}
        cls._db[key.lstrip("-")] = value

    @classmethod
    def clear(cls):
        cls._db = {}

    def hasName(self, key):
        return key in Options._db

    def getString(self, key, default=None):
        v = Options._db.get(key, default)
        return v if v is None or isinstance(v, bool) else str(v)

    def getInt(self, key, default=None):
        return int(Options._db[key]) if key in Options._db else default

    def getReal(self, key, default=None):
        return float(Options._db[key]) if key in Options._db else default

    def getBool(self, key, default=None):
        """PETSc spelling: a bare `-flag` is true; otherwise 1/0, true/false, yes/no, on/off."""
        if key not in Options._db:
            return default
        v = Options._db[key]
        if isinstance(v, bool):
            return v
        s = str(v).strip().lower()
        if s in ("1", "true", "yes", "on"):
            return True
        if s in ("0", "false", "no", "off"):
            return False
        raise Exception(f"option -{key}: cannot read '{v}' as a boolean")


def _is_number(s):
    try:
        float(s)
        return True
    except ValueError:
        return False
