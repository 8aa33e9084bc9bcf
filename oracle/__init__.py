"""CPU oracle for the KGE hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and there only as the checker or as the timed CPU
baseline -- never as the thing shipped.  The product package
(``customknowledgegraphembedding_b200``) never imports it and fails loudly when
its CUDA library is missing.

Parity status (see DESIGN.md "Oracle"):

* InterHT + the three modes + the fused TF loss terms + the TF train step are
  **pinned through the reference's own source**: ``oracle/run_reference_tf.py``
  executes ``/root/reference/tensorflow_codes/{model,supervisor}.py`` unmodified
  under ``oracle/tf_shim`` (a torch-backed stand-in for the ~25 TensorFlow
  symbols those files touch; TensorFlow itself is not installable here) and the
  outputs are committed under ``tests/golden/``.
* TransE / DistMult / ComplEx / RotatE / pRotatE, the PyTorch ``train_step`` and
  ``test_step`` live in the git submodule ``KnowledgeGraphEmbedding`` @
  2e440e0f9c687314d5ff67ead68ce985dc446e3a whose source is absent from
  ``/root/reference`` (``.SUBMODULES.json:8-16``).  They are restated here from
  the published algorithm: **parity
  unpinned** for those rows.
"""
