"""Host mirror of the reference's similarity-edge step.

    group = CompareGroup(sim_cfg)                  # J/compare/CompareGroup.java, Rdf2GrphConverter.java:47-55
    group.addToSource(v); group.addToTarget(v)     # while the triples are read (:99-110)
    vert, other, sim = group.compare(native, labels)        # the CompareJob loop (:127-186) — on the device
    edges = add_similarity_edges(edge_list, vert, other, sim)   # two directed edges per pair (:160-171)

What stays on the host is what the Java host would do in the shim: turning each literal's label into
the numbers the metric reads (shingle / token profiles as id lists, Integer.parseInt, LocalDate
fields).  The O(|source| x |target|) compare itself is ge_similarity_pairs in libgraphemb.so; there
is no CPU fallback.  JAROWINKLER and LEVENSHTEIN get the labels themselves (UTF-16 code units).
"""
from __future__ import annotations

import calendar
import ctypes as C
import re

import numpy as np

from . import _lib
from .synth import EdgeList

FLAG_EMPTY, FLAG_LONG, FLAG_PARSED = 1, 2, 4
NUM_OK, NUM_NFE, NUM_OOB = 0, 1, 2

_STOP = {"the", "of", "and", "a", "an", "to", "in", "is", "you", "that", "it", "for", "on", "from", "are", "as",
         "with", "at", "or", "by", "but", "if"}                       # TokenBased.java:13-18
_WS = re.compile("[ \t\n\x0b\f\r]+")                                   # java.util.regex \s+


def _units(s: str) -> str:
    """Java's view of a string: one element per UTF-16 code unit (only differs outside the BMP)."""
    if all(ord(c) < 0x10000 for c in s):
        return s
    b = s.encode("utf-16-le", "surrogatepass")
    return "".join(chr(int.from_bytes(b[k:k + 2], "little")) for k in range(0, len(b), 2))


def ngram_profile(label: str, k: int) -> dict:
    """info.debatty ShingleBased.getProfile: whitespace runs -> one blank, then k-shingle counts."""
    t = _WS.sub(" ", _units(label))
    prof = {}
    for i in range(len(t) - k + 1):
        sh = t[i:i + k]
        prof[sh] = prof.get(sh, 0) + 1
    return prof


def _trim(s: str) -> str:
    a, b = 0, len(s)
    while a < b and s[a] <= " ":
        a += 1
    while b > a and s[b - 1] <= " ":
        b -= 1
    return s[a:b]


def token_profile(label: str) -> dict:
    """TokenBased.getProfile (TokenBased.java:20-29; Tokenator :33-73): a token ends at a blank or
    at the last character, is trimmed, and counts when longer than one character and no stop word."""
    s = _units(label)
    prof, start = {}, 0
    for pos, ch in enumerate(s):
        if ch == " " or pos == len(s) - 1:
            tok = _trim(s[start:pos + 1])
            start = pos + 1
            if len(tok) > 1 and tok not in _STOP:
                prof[tok] = prof.get(tok, 0) + 1
    return prof


def parse_int(s: str):
    """Integer.parseInt (ASCII digits; None = NumberFormatException)."""
    if not s:
        return None
    body = s[1:] if s[0] in "+-" else s
    if not body or any(c < "0" or c > "9" for c in body):
        return None
    v = int(body)
    if s[0] == "-":
        v = -v
    return v if -2147483648 <= v <= 2147483647 else None


_OFFSET = re.compile(r"^(?:[Zz]|[+-](\d\d)(?:(\d\d)(\d\d)?)?)?$")


def _pattern_regex(pattern: str):
    """DateTimeFormatter.ofPattern for the letters a date needs: y/u (year), M (month), d (day);
    quoted text and other characters are literals.  Returns (compiled regex, group order)."""
    out, order, i = "", [], 0
    while i < len(pattern):
        c = pattern[i]
        if c == "'":
            j = pattern.find("'", i + 1)
            if j < 0:
                raise ValueError("unterminated quote in date pattern")
            out += re.escape(pattern[i + 1:j] or "'")
            i = j + 1
            continue
        if c in "yuMd":
            j = i
            while j < len(pattern) and pattern[j] == c:
                j += 1
            n = j - i
            if c in "yu":
                out += r"(\d{2})" if n == 2 else r"(\d{%d,19})" % n      # appendValue(field, n, 19, ...)
                order.append("y2" if n == 2 else "y")
            else:
                if n > 2:
                    raise ValueError("text month/day patterns are not supported: " + pattern)
                out += r"(\d{2})" if n == 2 else r"(\d{1,2})"
                order.append(c)
            i = j
            continue
        if c.isalpha():
            raise ValueError("unsupported date pattern letter %r in %r" % (c, pattern))
        out += re.escape(c)
        i += 1
    return re.compile("^" + out + "$"), order


def parse_local_date(text: str, pattern: str = "iso"):
    """LocalDate.parse(text, format) -> (year, month, day) or None (DateTimeParseException).
    "iso" = BASIC_ISO_DATE (yyyyMMdd + optional offset, ResolverStyle.STRICT); anything else goes
    through ofPattern's default SMART resolver (day 29-31 is clamped to the month's length)."""
    if pattern == "iso":
        if len(text) < 8 or not all("0" <= c <= "9" for c in text[:8]):
            return None
        y, m, d = int(text[:4]), int(text[4:6]), int(text[6:8])
        mo = _OFFSET.match(text[8:])
        if mo is None or not 1 <= m <= 12 or not 1 <= d <= _month_len(y, m):
            return None
        if mo.group(1) is not None:
            hh, mm, ss = int(mo.group(1)), int(mo.group(2) or 0), int(mo.group(3) or 0)
            if hh > 18 or mm > 59 or ss > 59 or (hh == 18 and (mm or ss)):
                return None
        return y, m, d
    rx, order = _pattern_regex(pattern)
    mo = rx.match(text)
    if mo is None or set(order) - {"y", "y2", "M", "d"} or not {"M", "d"} <= set(order) or not ({"y", "y2"} & set(order)):
        return None
    f = {}
    for name, g in zip(order, mo.groups()):
        v = int(g)
        if name == "y2":
            name, v = "y", 2000 + v                     # appendValueReduced(YEAR_OF_ERA, 2, 2, 2000)
        if f.setdefault(name, v) != v:
            return None
    y, m, d = f["y"], f["M"], f["d"]
    if not 1 <= m <= 12 or not 1 <= d <= 31 or y < 1:               # year-of-era starts at 1
        return None
    return y, m, min(d, _month_len(y, m))               # SMART


def _month_len(y: int, m: int) -> int:
    return 29 if m == 2 and calendar.isleap(y) else (31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31)[m - 1]


def epoch_day(y: int, m: int, d: int) -> int:
    """LocalDate.toEpochDay (proleptic ISO).  Python's date starts at year 1; the Gregorian
    calendar repeats every 400 years = 146 097 days, which carries year 0 and years past 9999."""
    import datetime
    shift = 0
    while y < 1:
        y, shift = y + 400, shift - 146097
    while y > 9999:
        y, shift = y - 400, shift + 146097
    return datetime.date(y, m, d).toordinal() - 719163 + shift


def java_hashset_order(ints) -> list:
    """Iteration order of a java.util.HashSet<Integer> after adding `ints` in order (table doubles
    from 16 when size exceeds 0.75 x capacity; buckets in index order, insertion order inside one —
    treeified buckets are not modelled).  This is the order of group.source.stream().toArray()
    (Rdf2GrphConverter.java:139-140)."""
    seen, uniq = set(), []
    for v in ints:
        if v not in seen:
            seen.add(v)
            uniq.append(int(v))
    cap = 16
    while len(uniq) > 0.75 * cap:
        cap *= 2
    def bucket(v):
        h = v & 0xFFFFFFFF
        return (h ^ (h >> 16)) & (cap - 1)
    return [v for _, _, v in sorted((bucket(v), k, v) for k, v in enumerate(uniq))]


class CompareGroup:
    """J/compare/CompareGroup.java + the metric of Configuration.SimilarityGroup.toFunction
    (Configuration.java:201-228).  `sim` is one entry of the config's `similarity` list."""

    def __init__(self, sim: dict):
        self.method = str(sim["method"]).upper()
        if self.method not in _lib.SIM_METHOD:
            raise ValueError("Unsupported similarity method on the device path: " + self.method)
        src = sim.get("sourcePredicate", sim.get("predicate"))
        dst = sim.get("targetPredicate", sim.get("predicate"))
        self.sourcePredicate, self.targetPredicate = src, dst
        self.upperTriangle = src == dst                                  # Rdf2GrphConverter.java:50
        self.threshold = float(sim.get("threshold", 0.0))
        self.ngram = int(sim.get("ngram") or 3)                          # getNgram
        self.smooth = float(sim.get("smooth") or 1.0)                    # getSmooth
        self.distance = float(sim.get("distance") or 0.0)
        self.pattern = sim.get("pattern") or "iso"
        self.time = str(sim.get("time") or "bidirectional").upper()
        if self.time not in _lib.SIM_TIME:
            raise ValueError("bad time direction: " + self.time)
        self._source, self._target = [], []

    def addToSource(self, i: int): self._source.append(int(i))
    def addToTarget(self, i: int): self._target.append(int(i))

    @property
    def source(self): return java_hashset_order(self._source)
    @property
    def target(self): return java_hashset_order(self._target)

    # ---- label -> numbers --------------------------------------------------------------------
    def _describe(self, labels: list):
        n = len(labels)
        units = [_units(s) for s in labels]
        flags = np.zeros(n, np.uint8)
        ids = {}
        sid = np.array([ids.setdefault(u, len(ids)) for u in units], np.int32)
        for k, u in enumerate(units):
            if not u:
                flags[k] |= FLAG_EMPTY
        keep = {"flags": flags, "sid": sid}
        m = self.method
        if m in ("NGRAM_COSINE", "NGRAM_JACCARD", "TOKEN_COSINE", "TOKEN_JACCARD"):
            vocab, off, keys, cnts = {}, np.zeros(n + 1, np.int64), [], []
            for k, u in enumerate(units):
                prof = ngram_profile(u, self.ngram) if m.startswith("NGRAM") else token_profile(u)
                if len(u) >= self.ngram:
                    flags[k] |= FLAG_LONG
                row = sorted((vocab.setdefault(sh, len(vocab)), c) for sh, c in prof.items())
                keys += [a for a, _ in row]
                cnts += [c for _, c in row]
                off[k + 1] = len(keys)
            keep.update(prof_off=off, prof_key=np.array(keys or [0], np.int32), prof_cnt=np.array(cnts or [0], np.int32))
        elif m == "NUMERIC":
            hats = [u.find("^") for u in units]
            cuts = sorted(set(hats))
            slot = {h: c for c, h in enumerate(cuts)}
            val = np.zeros((n, len(cuts)), np.int32)
            st = np.zeros((n, len(cuts)), np.uint8)
            for k, u in enumerate(units):
                for h, c in slot.items():
                    if h > len(u):
                        st[k, c] = NUM_OOB
                        continue
                    v = parse_int(u if h < 0 else u[:h])
                    if v is None:
                        st[k, c] = NUM_NFE
                    else:
                        val[k, c] = v
            keep.update(n_cut=len(cuts), cut_slot=np.array([slot[h] for h in hats], np.int32), num_val=val, num_status=st)
        elif m in ("JAROWINKLER", "LEVENSHTEIN"):
            codes = [np.frombuffer(u.encode("utf-16-le", "surrogatepass"), np.uint16) for u in units]
            off = np.zeros(n + 1, np.int64)
            np.cumsum([len(c) for c in codes], out=off[1:])
            keep.update(char_off=off, chars=np.concatenate(codes + [np.zeros(1, np.uint16)]))
        else:
            ep, pk = np.zeros(n, np.int64), np.zeros(n, np.int64)
            for k, u in enumerate(units):
                h = u.find("^")
                ymd = parse_local_date(u if h < 0 else u[:h], self.pattern)
                if ymd is not None:
                    flags[k] |= FLAG_PARSED
                    ep[k] = epoch_day(*ymd)
                    pk[k] = (ymd[0] * 12 + ymd[1] - 1) * 32 + ymd[2]
            keep.update(epoch_day=ep, packed_month=pk)
        return keep

    def compare(self, native, labels: dict):
        """labels: vertex id -> label string for every vertex of the group.  Returns (vert, other,
        similarity f32) in (source index, target index) order, and leaves `seconds` on self."""
        src, dst = self.source, self.target
        verts = sorted(set(src) | set(dst))
        item = {v: k for k, v in enumerate(verts)}
        d = self._describe([labels[v] for v in verts])
        vert = np.array(verts or [0], np.int32)
        s_idx = np.array([item[v] for v in src] or [0], np.int32)
        t_idx = np.array([item[v] for v in dst] or [0], np.int32)
        g = _lib.GeSimGroup()
        g.method, g.upper_triangle = _lib.SIM_METHOD[self.method], int(self.upperTriangle)
        g.threshold, g.smooth, g.distance = self.threshold, self.smooth, self.distance
        g.time_dir, g.n_items = _lib.SIM_TIME[self.time], len(verts)
        g.n_source, g.n_target, g.n_cut = len(src), len(dst), int(d.get("n_cut", 0))
        arrays = {"vert": vert, "string_id": d["sid"], "flags": d["flags"], "source": s_idx, "target": t_idx}
        for name in ("prof_off", "prof_key", "prof_cnt", "cut_slot", "num_val", "num_status", "epoch_day", "packed_month",
                     "char_off", "chars"):
            if name in d:
                arrays[name] = d[name]
        arrays = {k: np.ascontiguousarray(a) for k, a in arrays.items()}
        for k, a in arrays.items():
            setattr(g, k, a.ctypes.data_as(C.c_void_p))
        count, secs = C.c_int64(0), C.c_double(0.0)
        native.ck(native.L.ge_similarity_pairs(native.h, C.byref(g), C.byref(count), C.byref(secs)))
        self.seconds = secs.value
        n = count.value
        v, o, s = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float32)
        if n:
            native.ck(native.L.ge_similarity_export(native.h, n, v.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p),
                                                    s.ctypes.data_as(C.c_void_p)))
        return v, o, s


def add_similarity_edges(edges: EdgeList, vert, other, sim) -> EdgeList:
    """Rdf2GrphConverter.java:160-171: every kept pair becomes two directed edges (vert -> other,
    other -> vert) of type 0 whose weight is the similarity.  addDirectedSimpleEdge appends, so the
    new edges come after the triple edges, in pair order."""
    vert, other = np.asarray(vert, np.int32), np.asarray(other, np.int32)
    s = np.empty(2 * len(vert), np.int32); s[0::2], s[1::2] = vert, other
    t = np.empty(2 * len(vert), np.int32); t[0::2], t[1::2] = other, vert
    w = np.repeat(np.asarray(sim, np.float32), 2)
    return EdgeList(edges.V, np.concatenate([np.asarray(edges.src, np.int32), s]),
                    np.concatenate([np.asarray(edges.dst, np.int32), t]),
                    np.concatenate([np.asarray(edges.w, np.float32), w]))
