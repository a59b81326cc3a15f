"""f1, live: the product's metadata builder fed by the REFERENCE's real host objects, in process.
Runs only where /root/reference exists (the build container); the recorded twin of this test,
tests/test_host.py::test_metadata_builder_on_reference_host_trace, travels."""
import os

import pytest

pytestmark = pytest.mark.skipif(not os.path.isdir("/root/reference/vllm"),
                                reason="needs the reference checkout at /root/reference")


@pytest.mark.parametrize("method,bs,budget,lens", [("streamingLLM", 16, 128, [200, 131, 100]),
                                                   ("global", 16, 128, [300, 120])])
def test_product_builder_on_live_reference_objects(method, bs, budget, lens):
    """Every step the fork's Scheduler.schedule() -> ModelInputForGPUBuilder.add_seq_group() runs for
    real; the SAME builder object (its live InterDataForSeqGroup list, its runner proxy) is handed to
    this package's TorchCUDASDPAMetadataBuilder, whose metadata must equal what the fork's own
    attention metadata builder made of it -- through prefill, 45 decode steps and the
    update_block_tables feedback (scheduler.py:1338-1432,1879-1890, model_runner.py:530-546)."""
    from tests.ref_host_driver import RefHost, metadata_record
    from tests.test_host import _assert_md_equal
    from vllm_paged_compression_b200.backend import TorchCUDASDPAMetadataBuilder
    h = RefHost(method, budget, bs, lens)
    n_evict = 0
    for _ in range(46):
        b, seq_lens, query_lens, md_ref = h.step()
        # the fork's runner carries the fork's CacheConfig: the product builder reads
        # runner.cache_config.paged_evict_config (not None) and runner.device / pin_memory only
        md = TorchCUDASDPAMetadataBuilder(b).build(seq_lens, query_lens, -1, len(seq_lens))
        _assert_md_equal(md, metadata_record(md_ref))
        n_evict += len(h.finish_step(md_ref))
    assert n_evict > 0


def test_integration_patch_applies_to_the_fork():
    """integration/vllm_fork.patch (selector + _custom_ops edits of INTEGRATION.md, Options A and B)
    applies cleanly to the reference checkout (dry run: /root/reference is read-only)."""
    import subprocess
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(os.path.join(repo, "integration", "vllm_fork.patch")) as f:
        r = subprocess.run(["patch", "-p1", "--dry-run", "-d", "/root/reference"], stdin=f,
                           stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    assert "selector.py" in r.stdout and "_custom_ops.py" in r.stdout
