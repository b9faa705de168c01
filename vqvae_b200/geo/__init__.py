"""Geometric analysis and graph-based utilities (reference: src/geo/__init__.py:5-8)."""
from .knn_graph_optimized import build_knn_graph
from .geo_shortest_paths import dijkstra_multi_source

__all__ = ["build_knn_graph", "dijkstra_multi_source"]
