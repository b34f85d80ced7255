"""NAND-only netlists for BinFHEContext.EvalCircuit (mkfhe_eval_circuit).  The reference's multi-key paths implement the NAND
gate only (binfhe-base-scheme.cpp:341-342, 397), so every circuit a caller runs on them is a NAND netlist; these builders
produce the (in1, in2, out) rows EvalCircuit takes, in topological order, never with one wire on both inputs of a gate
(which the reference rejects, binfhe-base-scheme.cpp:382-383)."""
import numpy as np


class Netlist:
    """Grows a wire table: wires [0, n_inputs) are primary inputs, every nand() allocates the next wire."""

    def __init__(self, n_inputs: int):
        self.n_wires = int(n_inputs)
        self.gates = []

    def nand(self, x: int, y: int) -> int:
        if x == y:
            raise ValueError("a gate needs two different wires")
        self.gates.append((x, y, self.n_wires))
        self.n_wires += 1
        return self.n_wires - 1

    def xor(self, x: int, y: int):
        """(x xor y, NAND(x, y)) in 4 gates / 3 levels"""
        n = self.nand(x, y)
        return self.nand(self.nand(x, n), self.nand(y, n)), n

    def full_adder(self, a: int, b: int, c: int):
        """(sum, carry) in 9 gates; the carry leaves 2 levels after c arrives"""
        x, n1 = self.xor(a, b)
        s, n4 = self.xor(x, c)
        return s, self.nand(n1, n4)  # (a and b) or ((a xor b) and c)

    def array(self):
        return np.array(self.gates, dtype=np.uint32).reshape(-1, 3)


def ripple_carry_adders(bits: int, instances: int):
    """`instances` independent adders of two `bits`-bit numbers.  Primary inputs of adder t: wires [t * (2 bits + 1) ..):
    a_0..a_{bits-1}, b_0..b_{bits-1}, carry-in.  Returns (Netlist, sums) with sums[t] = the bits + 1 wires of a + b + carry-in,
    least significant first.  9 * bits gates per adder, 2 * bits + 4 levels."""
    per = 2 * bits + 1
    nl = Netlist(per * instances)
    sums = []
    for t in range(instances):
        c = per * t + 2 * bits
        out = []
        for i in range(bits):
            s, c = nl.full_adder(per * t + i, per * t + bits + i, c)
            out.append(s)
        sums.append(out + [c])
    return nl, sums
