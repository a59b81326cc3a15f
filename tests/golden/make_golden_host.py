#!/usr/bin/env python
"""Generate tests/golden/ref_host.json: the REFERENCE's real Scheduler / SelfAttnBlockSpaceManager /
ModelInputForGPUBuilder / TorchCUDASDPAMetadataBuilder driven on CPU (tests/ref_host_driver.py) for
prefill + decode steps with the eviction feedback loop.  Per step: the attributes of every
InterDataForSeqGroup the attention metadata builder reads, and the metadata the reference builder
produced from them.

    python tests/golden/make_golden_host.py        (needs /root/reference)
"""
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
from tests.ref_host_driver import RefHost, inter_record, metadata_record  # noqa: E402

RUNS = {
    "stream_b3": dict(evict_method="streamingLLM", budget=128, block_size=16, prompt_lens=[200, 131, 100], steps=70),
    "stream1_b2_bs32": dict(evict_method="streamingLLM-1", budget=256, block_size=32, prompt_lens=[1000, 257], steps=80),
    "global_b2": dict(evict_method="global", budget=128, block_size=16, prompt_lens=[128, 300], steps=40),
    "invkey_b2": dict(evict_method="inverse_key_l2", budget=128, block_size=16, prompt_lens=[150, 90], steps=50),
}


def main():
    out = {}
    for name, cfg in RUNS.items():
        steps = cfg.pop("steps")
        h = RefHost(**cfg)
        recs = []
        for _ in range(steps + 1):
            b, seq_lens, query_lens, md = h.step()
            rec = {"inter": [inter_record(d) for d in b.inter_data_list], "seq_lens": seq_lens,
                   "query_lens": query_lens, "md": metadata_record(md)}
            rmv = h.finish_step(md)
            rec["rmv"] = {str(r): list(v) for r, v in rmv.items()}
            recs.append(rec)
        out[name] = {"cfg": dict(cfg, steps=steps), "steps": recs}
        print(name, "evictions", sum(len(r["rmv"]) for r in recs))
    path = os.path.join(REPO, "tests", "golden", "ref_host.json")
    with open(path, "w") as f:
        json.dump(out, f, separators=(",", ":"))
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    main()
