"""nv_genai_b200 — B200-native exact dense retrieval behind LangChain's FAISS vector-store API.

Drop-in for the flat vector-store search path of april-yyt/NV_GenAI
(`FAISS(embedder, IndexFlatL2(d), InMemoryDocstore(), {})`,
/root/reference/RetrievalAugmentedGeneration/common/utils.py:217).  Importing this package
loads nv_genai_b200/lib/libb2flat.so (hand-written sm_100a CUDA behind the C ABI in
include/b2_flat.h); there is no CPU or PyTorch fallback.
"""
from . import _native
from .batching import BatchingIndex
from .documents import DistanceStrategy, Document, Embeddings, InMemoryDocstore
from .index import (METRIC_INNER_PRODUCT, METRIC_L2, IndexFlat, IndexFlatIP, IndexFlatL2, merge_shard_lists,
                    normalize_L2, synth_fill)
from .io import read_index, write_index
from .sharded import ShardedFlatIndex, ShardGroup, shard_bounds
from .vectorstore import FAISS, VectorStoreRetriever

B200FlatStore = FAISS  # the working name used in SURVEY.md 8b

__all__ = [
    "FAISS", "B200FlatStore", "VectorStoreRetriever", "IndexFlat", "IndexFlatL2", "IndexFlatIP",
    "METRIC_INNER_PRODUCT", "METRIC_L2", "normalize_L2", "read_index", "write_index", "Document", "Embeddings",
    "InMemoryDocstore", "DistanceStrategy", "ShardedFlatIndex", "ShardGroup", "BatchingIndex", "shard_bounds", "merge_shard_lists", "synth_fill",
]
__version__ = "0.1.0"
