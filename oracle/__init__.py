"""CPU oracle for the CTR scoring hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under this directory is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import it, and there only as the checker (or as the timed CPU baseline),
never as the thing shipped.  The product path (``distributed-tf-serving_b200/``)
must not import from here and fails loudly when its CUDA library is missing.

PARITY UNPINNED.  The reference (neuzxy/Distributed-TF-Serving) holds no tests,
no golden vectors and no recorded responses, and its model arithmetic lives in an
external TF-Serving process plus a SavedModel that the repository does not
contain.  What *is* pinned by the reference and restated
here bit-exactly:

* the split rule                  DCNClient.java:46-55   -> ``dcn_client.partition_list``
* the demo inputs                 DCNClient.java:57-74   -> ``dcn_client.get_feature_ids/wts``
* the tensor contract             DCNClient.java:76-81, 97-108
* ordered fan-out / merge         DCNClient.java:146-164 -> ``dcn_client.process_a_request``
* the ascending sort              DCNClient.java:195     -> ``dcn_client.java_float_sort``
* the wire encoding               tensor.proto:48,64; predict.proto:12-40 -> ``wire``

The model math (``model.py``) is builder-defined; its
known-answer anchors are closed-form LR cases and the "all candidates identical"
property of the reference's own demo inputs.
"""
