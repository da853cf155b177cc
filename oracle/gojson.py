"""Shared helpers for the oracle (test infrastructure; see oracle/__init__.py).

Restates the few Go stdlib / apimachinery behaviours the hot path leans on
(SURVEY.md Appendix B).  None of that code is under /root/reference (modules
are not vendored), so each helper names the call site it serves.
"""
import copy
import json

# ---------------------------------------------------------------- canonical JSON


def canon(tree):
    """Canonical bytes of a JSON tree: sorted keys, compact, UTF-8.

    This is the parity form: two generated objects are
    "identical" when their canonical bytes are identical.
    """
    return json.dumps(tree, sort_keys=True, separators=(",", ":"),
                      ensure_ascii=False).encode("utf-8")


def compact(tree):
    """Compact JSON bytes, insertion order kept (what a Go-written API object
    looks like on the wire, minus Go's own field order)."""
    return json.dumps(tree, separators=(",", ":"), ensure_ascii=False).encode("utf-8")


def deepcopy(tree):
    return copy.deepcopy(tree)


def nil(v):
    """Go nil slice/map/pointer <-> absent or JSON null."""
    return v is None


def get_map(obj, key):
    """obj[key] as a map for reading; Go reads of a nil map give zero values."""
    v = obj.get(key) if obj else None
    return v if isinstance(v, dict) else {}


def meta(obj):
    return obj.get("metadata") or {}


def has_annotation(m, key):
    """metav1.HasAnnotation: key presence only."""
    return key in get_map(m, "annotations")


# ---------------------------------------------------------------- strconv.ParseBool

_TRUE = {"1", "t", "T", "TRUE", "true", "True"}
_FALSE = {"0", "f", "F", "FALSE", "false", "False"}


def parse_bool(s):
    """strconv.ParseBool; returns (value, ok).  Call sites:
    odh-notebook-controller/controllers/notebook_controller.go:88,99."""
    if s in _TRUE:
        return True, True
    if s in _FALSE:
        return False, True
    return False, False


# ---------------------------------------------------------------- time.Parse(RFC3339)


def _days_from_civil(y, m, d):
    y -= m <= 2
    era = (y if y >= 0 else y - 399) // 400
    yoe = y - era * 400
    doy = (153 * (m + (-3 if m > 2 else 9)) + 2) // 5 + d - 1
    doe = yoe * 365 + yoe // 4 - yoe // 100 + doy
    return era * 146097 + doe - 719468


def _civil_from_days(z):
    z += 719468
    era = (z if z >= 0 else z - 146096) // 146097
    doe = z - era * 146097
    yoe = (doe - doe // 1460 + doe // 36524 - doe // 146096) // 365
    y = yoe + era * 400
    doy = doe - (365 * yoe + yoe // 4 - yoe // 100)
    mp = (5 * doy + 2) // 153
    d = doy - (153 * mp + 2) // 5 + 1
    m = mp + (3 if mp < 10 else -9)
    return y + (m <= 2), m, d


_MDAYS = [31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31]


def parse_rfc3339(s):
    """time.Parse(time.RFC3339, s) -> (unix_ns, utc_offset_seconds) or None.

    Layout 2006-01-02T15:04:05Z07:00.  Go accepts a fractional second after the
    seconds field even though the layout has none, requires the literal 'T',
    and requires 'Z' or +hh:mm / -hh:mm.  Pinned by
    notebook-controller/controllers/culling_controller_test.go:187,197,207.
    """
    if not isinstance(s, str):
        return None
    n = len(s)

    def num(i, w):
        if i + w > n:
            return None
        t = s[i:i + w]
        if not (t.isascii() and t.isdigit()):
            return None
        return int(t)

    if n < 20:
        return None
    y, mo, d = num(0, 4), num(5, 2), num(8, 2)
    h, mi, sec = num(11, 2), num(14, 2), num(17, 2)
    if None in (y, mo, d, h, mi, sec):
        return None
    if s[4] != "-" or s[7] != "-" or s[10] != "T" or s[13] != ":" or s[16] != ":":
        return None
    if not (1 <= mo <= 12) or h > 23 or mi > 59 or sec > 59:
        return None
    dim = _MDAYS[mo - 1] + (1 if mo == 2 and (y % 4 == 0 and (y % 100 != 0 or y % 400 == 0)) else 0)
    if not (1 <= d <= dim):
        return None
    i = 19
    ns = 0
    if i < n and s[i] in ".,":
        j = i + 1
        while j < n and s[j].isascii() and s[j].isdigit():
            j += 1
        if j == i + 1:
            return None
        frac = s[i + 1:j][:9]
        ns = int(frac.ljust(9, "0"))
        i = j
    if i >= n:
        return None
    if s[i] == "Z":
        if i + 1 != n:
            return None
        off = 0
    elif s[i] in "+-":
        if i + 6 != n or s[i + 3] != ":":
            return None
        oh, om = num(i + 1, 2), num(i + 4, 2)
        if oh is None or om is None or oh > 23 or om > 59:
            return None
        off = (oh * 3600 + om * 60) * (1 if s[i] == "+" else -1)
    else:
        return None
    secs = _days_from_civil(y, mo, d) * 86400 + h * 3600 + mi * 60 + sec - off
    return secs * 1_000_000_000 + ns, off


def format_rfc3339(unix_ns, off=0):
    """t.Format(time.RFC3339): whole seconds, 'Z' for UTC else +hh:mm."""
    secs = unix_ns // 1_000_000_000 + off
    days, rem = divmod(secs, 86400)
    y, m, d = _civil_from_days(days)
    h, rem = divmod(rem, 3600)
    mi, sec = divmod(rem, 60)
    if off == 0:
        z = "Z"
    else:
        a = abs(off)
        z = "%s%02d:%02d" % ("+" if off > 0 else "-", a // 3600, (a % 3600) // 60)
    return "%04d-%02d-%02dT%02d:%02d:%02d%s" % (y, m, d, h, mi, sec, z)


# Go's zero time.Time is 0001-01-01T00:00:00Z; time.Parse errors return it.
ZERO_TIME_NS = _days_from_civil(1, 1, 1) * 86400 * 1_000_000_000
