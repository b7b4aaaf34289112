"""`import faiss` -> b200knn, for running llm/rag.py with no edit at all: put this directory and
its parent on PYTHONPATH ahead of (or instead of) the faiss-cpu wheel.  Only the names rag.py
uses exist (llm/rag.py:61,62,71,74,80,82,86,116,153,186,190,191)."""
from b200knn import IndexFlatL2, METRIC_L2, read_index, write_index  # noqa: F401
