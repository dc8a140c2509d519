#!/usr/bin/env python
"""Short driver for ncu: the DBoW2 word lookup (V1) for 1 M descriptors against a k = 10, L = 6 tree."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import racing_slam_b200 as pkg  # noqa: E402
import torch  # noqa: E402

syn = pkg.synthetic
voc = syn.random_vocabulary(10, 6, 6)
q = syn.vocabulary_queries(voc, 1_000_000, 7)
with pkg.Context(0) as ctx:
    ctx.vocab_upload(0, voc["child_offset"], voc["child_ids"], voc["desc"], voc["weight"], voc["word_id"])
    dq = torch.from_numpy(q).cuda()
    dw = torch.empty(len(q), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 5):
        ctx.vocab_transform_dev(0, dq.data_ptr(), len(q), dw.data_ptr())
    ctx.sync()
    print("distinct words", len(np.unique(dw.cpu().numpy())))
