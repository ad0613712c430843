"""Stand-ins for the langchain_core / langchain_community types the FAISS vector store uses.

If langchain_core is importable its own classes are used, so objects interoperate with a
real LangChain stack; otherwise these minimal equivalents keep the same attribute names
(`Document.page_content/.metadata`, `InMemoryDocstore._dict/.add/.delete/.search`,
`DistanceStrategy`, `Embeddings`), which is what the reference's callers touch
(/root/reference/RetrievalAugmentedGeneration/common/utils.py:334-403 reads
`vectorstore.docstore._dict` and `doc.metadata`).
"""
from __future__ import annotations

import enum
from typing import Any, Dict, List, Optional, Union

try:  # pragma: no cover - langchain is not installed in the build image
    from langchain_core.documents import Document  # type: ignore
    from langchain_core.embeddings import Embeddings  # type: ignore

    HAVE_LANGCHAIN = True
except Exception:  # ModuleNotFoundError in this image
    HAVE_LANGCHAIN = False

    class Document:  # type: ignore[no-redef]
        """Text chunk plus metadata (langchain_core.documents.Document)."""

        type: str = "Document"

        def __init__(self, page_content: str, metadata: Optional[dict] = None, **kwargs: Any):
            self.page_content = page_content
            self.metadata = dict(metadata) if metadata is not None else {}
            for k, v in kwargs.items():
                setattr(self, k, v)

        def __setstate__(self, state: Any) -> None:
            """Also accepts what langchain's own Document pickles: langchain-core 0.1.x documents are pydantic-v1 models
            whose state is {'__dict__': fields, '__fields_set__': ..., '__private_attribute_values__': ...} (an
            `index.pkl` written by the reference's apps, e.g.
            /root/reference/experimental/oran-chatbot-multimodal/vectorstore/vectorstore_updater.py:220); pydantic-v2
            documents add '__pydantic_*__' keys around the same '__dict__'."""
            if isinstance(state, tuple) and len(state) == 2:  # (dict state, slots state)
                state = state[0] or {}
            if isinstance(state, dict) and isinstance(state.get("__dict__"), dict):
                state = state["__dict__"]
            self.__dict__.update(state)
            self.__dict__.setdefault("page_content", "")
            if self.__dict__.get("metadata") is None:
                self.__dict__["metadata"] = {}

        def __eq__(self, other: object) -> bool:
            return (isinstance(other, Document) and self.page_content == other.page_content
                    and self.metadata == other.metadata)

        def __hash__(self) -> int:  # documents are mutable; identity hash like pydantic v1 models lack
            return id(self)

        def __repr__(self) -> str:
            return f"Document(page_content={self.page_content!r}, metadata={self.metadata!r})"

    class Embeddings:  # type: ignore[no-redef]
        """Interface of langchain_core.embeddings.Embeddings."""

        def embed_documents(self, texts: List[str]) -> List[List[float]]:
            raise NotImplementedError

        def embed_query(self, text: str) -> List[float]:
            raise NotImplementedError


class DistanceStrategy(str, enum.Enum):
    """langchain_community.vectorstores.utils.DistanceStrategy."""

    EUCLIDEAN_DISTANCE = "EUCLIDEAN_DISTANCE"
    MAX_INNER_PRODUCT = "MAX_INNER_PRODUCT"
    DOT_PRODUCT = "DOT_PRODUCT"
    JACCARD = "JACCARD"
    COSINE = "COSINE"


class InMemoryDocstore:
    """langchain_community.docstore.in_memory.InMemoryDocstore: a dict id -> Document."""

    def __init__(self, _dict: Optional[Dict[str, Document]] = None):
        self._dict = _dict if _dict is not None else {}

    def add(self, texts: Dict[str, Document]) -> None:
        overlapping = set(texts).intersection(self._dict)
        if overlapping:
            raise ValueError(f"Tried to add ids that already exist: {overlapping}")
        self._dict = {**self._dict, **texts}

    def delete(self, ids: List[str]) -> None:
        overlapping = set(ids).intersection(self._dict)
        if not overlapping:
            raise ValueError(f"Tried to delete ids that does not  exist: {ids}")
        for _id in ids:
            self._dict.pop(_id)

    def search(self, search: str) -> Union[str, Document]:
        if search not in self._dict:
            return f"ID {search} not found."
        return self._dict[search]
