"""CPU oracle for the scoring + bootstrap hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package imports this; only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``
may, and there only as the checker or as the timed CPU baseline.

Parity status (see DESIGN.md "Oracle"):
  * scorer semantics   -- pinned against the reference's own ``get_yes_no_logprobs``
    (imported from /root/reference by tests/golden/make_golden.py; outputs committed under
    tests/golden/).  The transformer arithmetic itself lives in the third-party
    ``transformers`` package (unpinned ``>=4.30.0`` in the reference's requirements.txt:11;
    5.5.0 in this image), for which the reference holds no tests: that part is anchored on
    this image's HF CPU execution, "parity unpinned" by the reference itself.
  * encoder-decoder    -- ``t5_ref.py`` pinned against HF ``T5ForConditionalGeneration.generate`` and against goldens from the
    reference's own ``is_encoder_decoder`` branches (tests/golden/make_t5_golden.py, t5_golden.json).
  * LLM.int8           -- ``int8_ref.py`` restates bitsandbytes' published ``MatMul8bitLt``; the library is absent from
    /root/reference and from this image: **parity unpinned**.
  * Qwen-1 / Baichuan2 -- their math is the Llama block (+ q|k|v bias), pinned through HF's Qwen2 / Llama classes; the hub-side
    custom modelling code cannot be instantiated offline, so the checkpoint key maps are **parity unpinned**.
  * bootstrap          -- pinned against numpy's legacy RNG and against the reference functions'
    outputs on the shipped data.
"""
