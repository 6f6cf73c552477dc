"""Development aid: runs tests/test_gpu_similarity.py on the CPU with ge_similarity_pairs replaced
by a line-by-line Python emulation of csrc/sim.cu (same arrays, same rules).  Validates the host-side
label -> numbers code and the kernel's decision logic before GPU time is spent; not a test, not product."""
import math
import sys

sys.path.insert(0, ".")
import numpy as np

from graph_embeddings_b200 import similarity as S


def pair(g, d, a, b):
    fa, fb = int(d["flags"][a]), int(d["flags"][b])
    same = d["sid"][a] == d["sid"][b]
    m = g.method
    if m in ("NGRAM_COSINE", "NGRAM_JACCARD", "TOKEN_COSINE", "TOKEN_JACCARD"):
        if same:
            return 1.0, False
        if m == "NGRAM_COSINE" and not (fa & 2 and fb & 2):
            return 0.0, False
        off, key, cnt = d["prof_off"], d["prof_key"], d["prof_cnt"]
        ka = {int(key[p]): int(cnt[p]) for p in range(off[a], off[a + 1])}
        kb = {int(key[p]): int(cnt[p]) for p in range(off[b], off[b + 1])}
        assert list(key[off[a]:off[a + 1]]) == sorted(ka)
        inter = len(set(ka) & set(kb))
        if "JACCARD" in m:
            uni = len(ka) + len(kb) - inter
            return (inter / uni if uni else float("nan")), False
        dot = sum(ka[k] * kb[k] for k in ka if k in kb)
        na, nb = math.sqrt(sum(v * v for v in ka.values())), math.sqrt(sum(v * v for v in kb.values()))
        den = na * nb
        return (dot / den if den else float("nan")), False
    if m in ("JAROWINKLER", "LEVENSHTEIN"):
        if same:
            return 1.0, False
        off, ch = d["char_off"], d["chars"]
        s1, s2 = list(ch[off[a]:off[a + 1]]), list(ch[off[b]:off[b + 1]])
        n1, n2 = len(s1), len(s2)
        if m == "LEVENSHTEIN":
            mx = max(n1, n2)
            if not (1.0 - abs(n1 - n2) / mx >= g.threshold):
                return float("nan"), False
            if n1 == 0 or n2 == 0:
                dist = n1 + n2
            else:
                row = list(range(n2 + 1))
                for i in range(n1):
                    diag, left = i, i + 1
                    row[0] = left
                    for j in range(n2):
                        up = row[j + 1]
                        v = min(left + 1, up + 1, diag + (0 if s1[i] == s2[j] else 1))
                        row[j + 1] = v
                        diag, left = up, v
                dist = row[n2]
            return 1.0 - dist / mx, False
        big, small = (s1, s2) if n1 > n2 else (s2, s1)
        if not small:
            return 0.0, False
        rng_ = max(len(big) // 2 - 1, 0)
        fb_, fs_ = [False] * len(big), [False] * len(small)
        matches = 0
        for mi, c1 in enumerate(small):
            for xi in range(max(mi - rng_, 0), min(mi + rng_ + 1, len(big))):
                if not fb_[xi] and c1 == big[xi]:
                    fb_[xi] = fs_[mi] = True
                    matches += 1
                    break
        if matches == 0:
            return 0.0, False
        differ, xi = 0, 0
        for mi in range(len(small)):
            if not fs_[mi]:
                continue
            while not fb_[xi]:
                xi += 1
            differ += small[mi] != big[xi]
            xi += 1
        prefix = 0
        for k in range(len(small)):
            if s1[k] == s2[k]:
                prefix += 1
            else:
                break
        f = np.float32
        mm = f(matches)
        j = float((mm / f(n1) + mm / f(n2) + (mm - f(differ // 2)) / mm) / f(3))
        if not j > 0.7:
            return j, False
        return j + min(0.1, 1.0 / len(big)) * prefix * (1 - j), False
    if (fa | fb) & 1:
        return 0.0, False
    if same:
        return 1.0, False
    if m == "NUMERIC":
        slot = d["cut_slot"][a]
        sb = d["num_status"][b, slot]
        if sb == 2:
            return 0.0, True
        if d["num_status"][a, slot] or sb:
            return 0.0, False
        diff = (int(d["num_val"][a, slot]) - int(d["num_val"][b, slot])) & 0xFFFFFFFF
        if diff >= 2 ** 31:
            diff -= 2 ** 32
        ad = diff if diff >= 0 else ((-diff + 2 ** 31) % 2 ** 32) - 2 ** 31
        base = float(ad)
    else:
        if not (fa & 4 and fb & 4):
            return 0.0, False
        d1, d2 = int(d["epoch_day"][a]), int(d["epoch_day"][b])
        if g.time == "BACKWARDS" and d1 > d2:
            return 0.0, False
        if g.time == "FORWARDS" and d1 < d2:
            return 0.0, False
        if m == "DATE_DAYS":
            between = d2 - d1
        else:
            q = int(d["packed_month"][b]) - int(d["packed_month"][a])
            between = abs(q) // 32 * (1 if q >= 0 else -1)
            if m == "DATE_YEARS":
                between = abs(between) // 12 * (1 if between >= 0 else -1)
        base = abs(float(between))
    return math.pow(abs(base - g.distance) + 1, g.smooth - 1), False


def compare(self, native, labels):
    src, dst = self.source, self.target
    verts = sorted(set(src) | set(dst))
    item = {v: k for k, v in enumerate(verts)}
    d = self._describe([labels[v] for v in verts])
    V, O_, S_ = [], [], []
    for i, sv in enumerate(src):
        a = item[sv]
        row, dead = [], False
        for j in range(i + 1 if self.upperTriangle else 0, len(dst)):
            b = item[dst[j]]
            if verts[b] == verts[a]:
                continue
            sim, oob = pair(self, d, a, b)
            dead |= oob
            if not oob and sim >= self.threshold:
                row.append((verts[b], np.float32(sim)))
        if not dead:
            for o, s in row:
                V.append(verts[a]); O_.append(o); S_.append(s)
    self.seconds = 1.0
    return np.array(V, np.int32), np.array(O_, np.int32), np.array(S_, np.float32)


if __name__ == "__main__":
    S.CompareGroup.compare = compare
    sys.path.insert(0, "tests")
    import test_gpu_similarity as T
    for method, thr in [("NGRAM_JACCARD", 0.45), ("NGRAM_COSINE", 0.6), ("TOKEN_JACCARD", 0.3), ("TOKEN_COSINE", 0.45)]:
        for upper in (True, False):
            T.test_profile_metrics_match_oracle(None, method, thr, upper); print("ok", method, upper, flush=True)
    T.test_threshold_zero_keeps_everything_but_nan(None); print("ok thr0")
    for upper in (True, False):
        T.test_numeric_matches_oracle_including_the_cut_quirk(None, upper); print("ok numeric", upper, flush=True)
    for m in ("DATE_DAYS", "DATE_MONTHS", "DATE_YEARS"):
        for t in ("bidirectional", "backwards", "forwards"):
            T.test_dates_match_oracle(None, m, t); print("ok", m, t, flush=True)
    for m, thr in (("JAROWINKLER", 0.9), ("LEVENSHTEIN", 0.75)):
        for upper in (True, False):
            T.test_string_metrics_match_oracle(None, m, thr, upper); print("ok", m, upper, flush=True)
    if len(sys.argv) > 1:
        T.test_many_segments_and_scan_tiles(None); print("ok big")
