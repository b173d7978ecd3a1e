"""CPU oracle for the EventGPT event -> visual-token front end.

TEST INFRASTRUCTURE ONLY.  Nothing under ``eventgpt_b200/`` may import, call,
link or execute anything in this package; the only legal users are ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs
of ``bench.py`` -- and there only as the checker / timed CPU baseline, never as
the thing shipped.

Every function restates one piece of the reference (``/root/reference``,
xifan-zhang/EventGPT) and cites the file:line it follows.  The restatement is
pinned against the real reference code: ``oracle/make_golden.py`` imports the
reference in the build container, runs it on the reference's own fixtures and
on adversarial inputs, and commits the results under ``tests/golden/``;
``tests/test_oracle_golden.py`` then checks this package against those vectors
on any machine (the reference itself cannot travel to the GPU box).

Third-party arithmetic the reference delegates to (not vendored upstream):
Pillow 10.4.0 ``Resample.c`` (bicubic), transformers 4.37.2 ``image_processing_clip``
and ``modeling_clip``, torch 2.1.2 Linear/GELU/LayerNorm.  The goldens were
produced with the versions installed in the image (Pillow 12.2.0, transformers
5.5.0 with the PIL-backed ``CLIPImageProcessorPil``, torch 2.11.0), which
SURVEY.md section 0/F2 shows reproduce the pinned behaviour exactly.
"""
