"""CPU: the NAND netlist builders (mkfhe_b200/netlist.py) and the levelisation the library applies to them
(mkfhe_circuit_levels) -- host logic of the circuit entry points, no GPU."""
import numpy as np
import pytest

import mkfhe_b200 as mk
from mkfhe_b200.netlist import Netlist, ripple_carry_adders


def _simulate(n_wires, gates, inputs):
    w = np.zeros(n_wires, dtype=np.int64)
    w[:len(inputs)] = inputs
    for a, b, o in gates:
        w[o] = 1 - (w[a] & w[b])
    return w


@pytest.mark.parametrize("bits,inst", [(1, 1), (4, 3), (8, 5)])
def test_ripple_carry_adders_add(bits, inst):
    nl, sums = ripple_carry_adders(bits, inst)
    per = 2 * bits + 1
    assert len(nl.gates) == 9 * bits * inst and nl.n_wires == per * inst + 9 * bits * inst
    rng = np.random.default_rng(bits)
    a, b, c = rng.integers(0, 1 << bits, inst), rng.integers(0, 1 << bits, inst), rng.integers(0, 2, inst)
    inputs = []
    for t in range(inst):
        inputs += [(int(a[t]) >> i) & 1 for i in range(bits)] + [(int(b[t]) >> i) & 1 for i in range(bits)] + [int(c[t])]
    w = _simulate(nl.n_wires, nl.gates, inputs)
    for t in range(inst):
        assert sum(int(w[s]) << i for i, s in enumerate(sums[t])) == int(a[t]) + int(b[t]) + int(c[t])
    levels, lev = mk.CircuitLevels(nl.n_wires, nl.array())
    assert levels == 2 * bits + 4  # the carry chain: two levels per bit after the first sum
    assert all(x != y for x, y, _ in nl.gates)  # the reference rejects a gate on one ciphertext
    # levels are consistent: every gate sits one above its deepest producer
    produced = {int(o): int(l) for (_, _, o), l in zip(nl.gates, lev)}
    for (x, y, _), l in zip(nl.gates, lev):
        assert int(l) == max(produced.get(x, -1), produced.get(y, -1)) + 1


def test_xor_truth_table_and_guard():
    for a in (0, 1):
        for b in (0, 1):
            nl = Netlist(2)
            x, n = nl.xor(0, 1)
            w = _simulate(nl.n_wires, nl.gates, [a, b])
            assert w[x] == a ^ b and w[n] == 1 - (a & b)
    with pytest.raises(ValueError):
        Netlist(2).nand(1, 1)
