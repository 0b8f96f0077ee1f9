"""Stand-in for the ``looseversion`` package (dotted-version comparison), needed to import the
unmodified reference (``thunder/executors/torchex.py:9`` and others). Baseline infrastructure only."""

import re

_component_re = re.compile(r"(\d+|[a-z]+|\.)", re.VERBOSE)


class LooseVersion:
    def __init__(self, vstring=None):
        self.vstring = ""
        self.version = []
        if vstring is not None:
            self.parse(str(vstring))

    def parse(self, vstring):
        self.vstring = vstring
        parts = [p for p in _component_re.split(vstring) if p and p != "."]
        out = []
        for p in parts:
            try:
                out.append(int(p))
            except ValueError:
                out.append(p)
        self.version = out

    def __str__(self):
        return self.vstring

    def __repr__(self):
        return f"LooseVersion ('{self.vstring}')"

    def _cmp(self, other):
        if isinstance(other, str):
            other = LooseVersion(other)
        elif not isinstance(other, LooseVersion):
            return NotImplemented
        a, b = self.version, other.version
        for x, y in zip(a, b):
            if x == y:
                continue
            if isinstance(x, int) and isinstance(y, int) or isinstance(x, str) and isinstance(y, str):
                return -1 if x < y else 1
            # mixed int/str: ints sort after strings (a pre-release tag is older)
            return 1 if isinstance(x, int) else -1
        return (len(a) > len(b)) - (len(a) < len(b))

    def __eq__(self, o):
        c = self._cmp(o)
        return c if c is NotImplemented else c == 0

    def __lt__(self, o):
        c = self._cmp(o)
        return c if c is NotImplemented else c < 0

    def __le__(self, o):
        c = self._cmp(o)
        return c if c is NotImplemented else c <= 0

    def __gt__(self, o):
        c = self._cmp(o)
        return c if c is NotImplemented else c > 0

    def __ge__(self, o):
        c = self._cmp(o)
        return c if c is NotImplemented else c >= 0

    def __hash__(self):
        return hash(self.vstring)
