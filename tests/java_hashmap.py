"""java.util.HashMap<Integer, V> (OpenJDK 8, the reference's target: pom.xml:13-14) as far as ITERATION
ORDER goes — the one thing BCV.toUnity's f32 sum depends on (J/bca/util/BCV.java:64-70,96-98).
Restated from the published algorithm (no JVM here): table of power-of-two capacity starting at 16,
load factor 0.75, index = (h ^ h >>> 16) & (cap - 1) with h = Integer.hashCode = the value;
put() appends a new key at the TAIL of its bin and doubles the table when ++size > threshold;
merge() doubles FIRST when size > threshold, and links a new key at the HEAD of its bin; a doubling
splits every bin into a low and a high list that keep their relative order; remove() unlinks.
Iteration walks the table by index, each bin along its chain.  Bins of 9 or more nodes would be
treeified (or force a doubling below 64 buckets): doubling is reproduced, treeification is not —
the emulator raises, and the tests use graphs where it does not happen."""


class TreeifiedBin(NotImplementedError):
    pass


class JavaHashMap:
    def __init__(self):
        self.table = None          # list of bins; a bin is a list of [key, value] in chain order
        self.threshold = 0
        self.size = 0

    @staticmethod
    def _hash(key):
        h = key & 0xFFFFFFFF
        return h ^ (h >> 16)

    def _resize(self):
        if self.table is None:
            self.table, self.threshold = [[] for _ in range(16)], 12
            return
        old = self.table
        cap = len(old)
        new = [[] for _ in range(2 * cap)]
        for j, chain in enumerate(old):
            for node in chain:         # lo / hi split, relative order preserved
                new[j + cap if self._hash(node[0]) & cap else j].append(node)
        self.table, self.threshold = new, self.threshold * 2

    def _treeify(self):
        if len(self.table) < 64:       # treeifyBin: MIN_TREEIFY_CAPACITY
            self._resize()
        else:
            raise TreeifiedBin("a bin reached the treeify threshold at >= 64 buckets")

    def _bin(self, key):
        return self.table[self._hash(key) & (len(self.table) - 1)]

    def get(self, key, default=None):
        if self.table is None:
            return default
        for node in self._bin(key):
            if node[0] == key:
                return node[1]
        return default

    def put(self, key, value):
        if self.table is None:
            self._resize()
        chain = self._bin(key)
        for node in chain:
            if node[0] == key:
                node[1] = value
                return
        chain.append([key, value])     # tail
        if len(chain) - 1 >= 8:        # binCount >= TREEIFY_THRESHOLD - 1 when the tail was reached
            self._treeify()
        self.size += 1
        if self.size > self.threshold:
            self._resize()

    def merge(self, key, value, fn):
        if self.table is None or self.size > self.threshold:
            self._resize()
        chain = self._bin(key)
        for node in chain:
            if node[0] == key:
                node[1] = fn(node[1], value)
                return
        existing = len(chain)
        chain.insert(0, [key, value])  # newNode(hash, key, value, first): HEAD of the bin
        if existing >= 7:
            self._treeify()
        self.size += 1

    def remove(self, key):
        if self.table is None:
            return
        chain = self._bin(key)
        for k, node in enumerate(chain):
            if node[0] == key:
                del chain[k]
                self.size -= 1
                return

    def items(self):
        if self.table is None:
            return
        for chain in self.table:
            for key, value in chain:
                yield key, value

    def keys(self):
        return [k for k, _ in self.items()]
