"""Host slicer / planner robustness: truncated and corrupted IPC streams must end in a B2AError (IOException /
InvalidInputException / NotImplementedException like the reference's reader) or in a valid plan whose every buffer lies
inside the stream image — never in a crash, and never in a plan the kernels would read out of bounds with."""
import os

import numpy as np

from parity import GOLDEN

FILES = ["data/test.arrows", "data/fruit.arrow", "arrow-testing/1.0.0-littleendian/generated_nested.stream",
         "arrow-testing/1.0.0-littleendian/generated_dictionary.stream",
         "arrow-testing/2.0.0-compression/generated_zstd.stream"]


def test_mutated_streams_never_crash(b2a):
    rng = np.random.default_rng(7)
    planned = rejected = 0
    for f in FILES:
        data = open(os.path.join(GOLDEN, f), "rb").read()
        for it in range(90):
            b = bytearray(data)
            if it % 3 == 0:
                b = b[:int(rng.integers(0, len(b)))]
            elif it % 3 == 1:
                for _ in range(int(rng.integers(1, 6))):
                    lim = min(len(b), 4096) if rng.random() < 0.7 else len(b)
                    b[int(rng.integers(0, lim))] = int(rng.integers(0, 256))
            else:
                p = int(rng.integers(0, max(1, len(b) - 8)))
                b[p:p + 8] = rng.integers(0, 256, 8, dtype=np.uint8).tobytes()
            try:
                r = b2a.ArrowIPCReader.open_buffers([bytes(b)])
                s = b2a.ArrowScan(r, device=-1)
            except b2a.B2AError:
                rejected += 1
                continue
            planned += 1
            assert s.image_lo + s.image_bytes <= len(b) + 16  # upload ranges are rounded to 16 bytes
            for i in range(r.num_batches):
                bi = r.batch_info(i)
                assert bi.body_offset + bi.body_length <= len(b)
    assert planned > 0 and rejected > 0
