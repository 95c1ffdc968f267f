"""Utilities: the counterparts of the reference's vendored ``resnet/official/utils`` tree.

``flags``        typed flag groups (base / performance / image / benchmark) without absl globals
``logs``         benchmark loggers, training hooks, hook factory, cloud probe (no network)
``misc``         distribution-strategy selection, stop threshold, synthetic data, clean
``export``       inference bundle export (the SavedModel receiver's role)
``testing``      golden-data harness, TF-bundle reader, synthetic integration runner, mock logger
``data``         datasets + batch iteration; DataFrame -> sharded record buffers (``file_io``)
``accelerator``  TPU helpers of the reference are out of scope; ``embedding_matmul`` is kept
"""
