from vqvae_b200.geo import build_knn_graph, dijkstra_multi_source

__all__ = ["build_knn_graph", "dijkstra_multi_source"]
