"""java.util.Random restated for tests (known answers: tests/test_oracle.py)."""


class JavaRandom:
    """java.util.Random (48-bit LCG): seeded runs of the reference draw exactly these numbers."""

    MASK = (1 << 48) - 1

    def __init__(self, seed):
        self.seed = (seed ^ 0x5DEECE66D) & self.MASK

    def next(self, bits):
        self.seed = (self.seed * 0x5DEECE66D + 0xB) & self.MASK
        r = self.seed >> (48 - bits)
        if r >= 1 << (bits - 1):
            r -= 1 << bits
        return r

    def next_double(self):
        a = self.next(26) & ((1 << 26) - 1)
        b = self.next(27) & ((1 << 27) - 1)
        return ((a << 27) + b) * (1.0 / (1 << 53))

