"""Drives the REFERENCE's own host objects -- `Scheduler`, `SelfAttnBlockSpaceManager`,
`SequenceGroup`, `ModelInputForGPUBuilder` and its `TorchCUDASDPAMetadataBuilder` -- on CPU for a
prefill + decode-with-eviction run.  Needs /root/reference; used by
tests/golden/make_golden_host.py (fixture) and tests/test_ref_live.py (live, skipped elsewhere).

What runs is the fork's real control flow either side of the attention backend:
  scheduler.schedule()                     vllm/core/scheduler.py:1300-1432 (seq_kv_lens :1338-1361)
  ModelInputForGPUBuilder.add_seq_group    vllm/worker/model_runner.py:455-546,740-790
  attn_metadata_builder.build              vllm/attention/backends/torch_cuda_sdpa.py:461-697
  scheduler.update_block_tables(req, idx)  llm_engine.py:1179-1193 -> scheduler.py:1879-1890
The model itself is replaced by "append one token per running sequence" and the layer-0 eviction
decision by the streamingLLM rule (victim = table entry 1 on trigger steps), which is what layer 0
returns for --evict-method streamingLLM (torch_cuda_sdpa.py:1404-1405); for local / global a fixed
pattern of middle-block indices stands in for the layer's choice.
"""
import sys
import types
import weakref

REF = "/root/reference"


def import_reference():
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import triton.runtime.cache as _trc
    for _n in ("default_cache_dir", "default_dump_dir", "default_override_dir"):
        if not hasattr(_trc, _n):
            setattr(_trc, _n, lambda: "/tmp/triton_stub")
    import vllm.attention.backends.torch_cuda_sdpa as HEAD
    import vllm.worker.model_runner as MR
    from vllm.config import CacheConfig, PagedEvictConfig, SchedulerConfig
    from vllm.core.scheduler import Scheduler
    from vllm.sampling_params import SamplingParams
    from vllm.sequence import Sequence, SequenceGroup
    return types.SimpleNamespace(HEAD=HEAD, MR=MR, CacheConfig=CacheConfig, PagedEvictConfig=PagedEvictConfig,
                                 SchedulerConfig=SchedulerConfig, Scheduler=Scheduler,
                                 SamplingParams=SamplingParams, Sequence=Sequence, SequenceGroup=SequenceGroup)


INTER_FIELDS = ["request_id", "is_prompt", "seq_ids", "block_tables", "orig_seq_lens", "seq_lens", "query_lens",
                "context_lens", "seq_kv_lens", "prefix_cache_hit"]
MD_LISTS = ["seq_lens", "seq_kv_lens"]
MD_INTS = ["num_prefills", "num_prefill_tokens", "num_decode_tokens", "max_query_len", "max_prefill_seq_len",
           "max_decode_seq_len"]
MD_TENSORS = ["slot_mapping", "seq_lens_tensor", "seq_kv_lens_tensor", "block_tables", "query_start_loc",
              "seq_start_loc", "context_lens_tensor"]


def metadata_record(md):
    rec = {k: list(getattr(md, k)) if getattr(md, k) is not None else None for k in MD_LISTS}
    rec.update({k: int(getattr(md, k)) for k in MD_INTS})
    rec.update({k: getattr(md, k).cpu().tolist() for k in MD_TENSORS})
    rec["reqid_mapping_blocktables"] = {r: {str(s): list(t) for s, t in d.items()}
                                        for r, d in md.reqid_mapping_blocktables.items()}
    return rec


def inter_record(d):
    rec = {}
    for f in INTER_FIELDS:
        v = getattr(d, f)
        if isinstance(v, dict):
            v = {str(k): (list(x) if isinstance(x, (list, tuple)) else x) for k, x in v.items()}
        rec[f] = v
    rec["input_token_lens"] = [len(t) for t in d.input_tokens]
    return rec


class _Runner:
    """Attributes ModelInputForGPUBuilder and the metadata builder read from GPUModelRunnerBase."""

    def __init__(self, **kw):
        self.__dict__.update(kw)


class RefHost:
    """The reference's scheduler + model-input builder for a fixed set of prompts."""

    def __init__(self, evict_method, budget, block_size, prompt_lens, num_gpu_blocks=512, topk=3):
        R = self.R = import_reference()
        self.bs, self.budget, self.method = block_size, budget, evict_method
        pe = R.PagedEvictConfig(False, "budget", evict_method, budget, topk)
        self.cache_config = cc = R.CacheConfig(block_size, 0.9, 0, "auto", paged_evict_config=pe)
        cc.num_gpu_blocks, cc.num_cpu_blocks = num_gpu_blocks, 0
        max_len = max(prompt_lens) + 4096
        self.scheduler_config = sc = R.SchedulerConfig(
            runner_type="generate", max_num_batched_tokens=max(sum(prompt_lens), max_len), max_num_seqs=256,
            max_model_len=max_len, enable_chunked_prefill=False)
        self.scheduler = R.Scheduler(sc, cc, None)
        self.runner = _Runner(
            inter_data_cache={}, _model_input_cls=R.MR.ModelInputForGPUWithSamplingMetadata,
            attn_backend=R.HEAD.TorchCUDASDPABackend, scheduler_config=sc, sliding_window=None,
            block_size=block_size, lora_config=None, prompt_adapter_config=None, multi_modal_input_mapper=None,
            cache_config=cc, device="cpu", pin_memory=False,
            model_config=types.SimpleNamespace(is_encoder_decoder=False, uses_mrope=False, enforce_eager=True),
            mm_registry=types.SimpleNamespace(has_processor=lambda mc: False))
        self.seqs = []
        for i, n in enumerate(prompt_lens):
            seq = R.Sequence(i, {"prompt_token_ids": list(range(n)), "prompt": None, "type": "token"}, block_size)
            sg = R.SequenceGroup(f"r{i}", [seq], arrival_time=0.0, sampling_params=R.SamplingParams(max_tokens=10 ** 6))
            self.scheduler.add_seq_group(sg)
            self.seqs.append(seq)

    def step(self):
        """One engine step up to the attention metadata.  -> (input builder holding the step's
        inter_data_list, seq_lens, query_lens, reference metadata)."""
        sgml, outs, _ = self.scheduler.schedule()
        b = self.R.MR.ModelInputForGPUBuilder(weakref.proxy(self.runner))
        self._keep = b
        for sgm in sgml:
            b.add_seq_group(sgm)
        seq_lens, query_lens = [], []
        for d in b.inter_data_list:              # model_runner.py ModelInputForGPUBuilder.build()
            seq_lens.extend(d.seq_lens)
            query_lens.extend(d.query_lens)
        md = b.attn_metadata_builder.build(seq_lens, query_lens, -1, len(seq_lens))
        self._sgml = sgml
        return b, seq_lens, query_lens, md

    def finish_step(self, md):
        """The model ran: every scheduled sequence gets a token; layer 0's decision goes back to
        the scheduler.  -> {row: (request_id, idx)} applied."""
        rmv = {}
        self.n_steps = getattr(self, "n_steps", 0) + 1
        if md.num_decode_tokens > 0 and self.method != "inverse_key_l2":
            rows = [(rid, sid) for rid, tabs in md.reqid_mapping_blocktables.items() for sid in tabs]
            for row, sl in enumerate(md.seq_lens):
                if sl > self.budget and sl % self.bs == 0:
                    # local / global: which middle block layer 0 picks is the kernel's business; the
                    # host only needs SOME valid index -- a fixed pattern stands in for it
                    idx = 1 if self.method.startswith("streamingLLM") else 1 + (self.n_steps + row) % 3
                    rmv[row] = (rows[row][0], idx)
        for sgm in self._sgml:
            for sid, sd in sgm.seq_data.items():
                sd.update_num_computed_tokens(sgm.token_chunk_size if sgm.is_prompt else 1)
        for seq in self.seqs:
            seq.append_token_id(len(seq.get_token_ids()), {len(seq.get_token_ids()): self._lp()})
        for row, (rid, idx) in rmv.items():
            self.scheduler.update_block_tables(rid, idx)        # llm_engine.py:1179-1193
        return rmv

    def _lp(self):
        from vllm.sequence import Logprob
        return Logprob(0.0)
