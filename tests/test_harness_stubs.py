"""The networkx stand-in of the N4 acceptance harness (tests/stubs/networkx) against hand-computed
centralities (networkx's own normalisation)."""
import importlib.util
from pathlib import Path


def test_networkx_stub_centralities():
    spec = importlib.util.spec_from_file_location("nx_stub", Path(__file__).parent / "stubs" / "networkx" / "__init__.py")
    nx = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(nx)
    G = nx.Graph()
    for a, b in [(0, 1), (1, 2), (2, 3), (1, 3), (3, 4)]:
        G.add_edge(a, b, num_inliers=10 * a + b)
    assert nx.degree_centrality(G) == {0: 0.25, 1: 0.75, 2: 0.5, 3: 0.75, 4: 0.25}
    assert nx.betweenness_centrality(G) == {0: 0.0, 1: 0.5, 2: 0.0, 3: 0.5, 4: 0.0}
    assert G.has_edge(3, 1) and not G.has_edge(0, 4) and G[1][3]["num_inliers"] == 13
    assert sorted(G.neighbors(3)) == [1, 2, 4] and len(G) == 5
    star = nx.Graph()
    for k in range(1, 5):
        star.add_edge(0, k)
    assert nx.betweenness_centrality(star)[0] == 1.0
