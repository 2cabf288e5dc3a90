"""The 16-key sorting network inside the shared-memory tile sort (gsworld_b200/csrc/binsort.cu, sort16) is
checked exhaustively with the 0-1 principle: a comparator network sorts every input iff it sorts all 2^16
binary inputs.  The comparator list is parsed from the CUDA source, so the test follows the kernel."""
import os
import re

import numpy as np

SRC = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gsworld_b200", "csrc", "binsort.cu")


def _network():
    text = open(SRC).read()
    body = text[text.index("void sort16("):]
    body = body[: body.index("\n}\n")]
    return [(int(a), int(b)) for a, b in re.findall(r"cswap\(k\[(\d+)\], k\[(\d+)\]\)", body)]


def test_sort16_network_sorts_all_binary_inputs():
    net = _network()
    assert len(net) == 60 and all(0 <= a < b < 16 for a, b in net)
    x = np.arange(1 << 16, dtype=np.uint32)
    bits = ((x[:, None] >> np.arange(16, dtype=np.uint32)[None, :]) & 1).astype(np.uint8)  # [65536, 16]
    for a, b in net:
        lo = np.minimum(bits[:, a], bits[:, b])
        hi = np.maximum(bits[:, a], bits[:, b])
        bits[:, a], bits[:, b] = lo, hi
    assert (np.diff(bits.astype(np.int8), axis=1) >= 0).all()
