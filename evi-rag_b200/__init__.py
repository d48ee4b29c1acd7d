"""B200-native retrieval hot path for EVI-RAG: triple scoring, DDE/BFS graph features, ragged top-k.

Host side mirrors the reference's Python interfaces (``Retriever.forward(batch) -> RetrieverOutput``,
``_bfs_dist`` ..., per-question ``topk``); compute runs in hand-written sm_100a kernels behind the
C-ABI library ``libevi_rag_b200.so`` (declared in ``include/evi_rag_b200.h``). There is no CPU
fallback: importing the compute modules without the built library raises.
"""
__version__ = "0.1.0"
