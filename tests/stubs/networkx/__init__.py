"""Minimal stand-in for the part of networkx the reference's image selector uses
(utils/image_selector.py:30-113): an undirected attribute graph, degree and betweenness centrality
(Brandes, unweighted, normalised as networkx does).  TEST INFRASTRUCTURE for
tests/run_reference_reconstruct.py only -- networkx is not in this image; the layout / drawing
calls of visualize_graph are no-ops."""
from collections import deque


class Graph:
    def __init__(self):
        self._adj = {}

    def add_edge(self, u, v, **attr):
        self._adj.setdefault(u, {})
        self._adj.setdefault(v, {})
        d = self._adj[u].get(v, {})
        d.update(attr)
        self._adj[u][v] = d
        self._adj[v][u] = d

    def nodes(self):
        return list(self._adj)

    def neighbors(self, n):
        return iter(self._adj[n])

    def has_edge(self, u, v):
        return u in self._adj and v in self._adj[u]

    def edges(self, data=False):
        seen, out = set(), []
        for u, nb in self._adj.items():
            for v, d in nb.items():
                if (v, u) not in seen:
                    seen.add((u, v))
                    out.append((u, v, d) if data else (u, v))
        return out

    def subgraph(self, nodes):
        keep = set(nodes)
        g = Graph()
        for u in self._adj:
            if u in keep:
                g._adj.setdefault(u, {})
        for u, v, d in self.edges(data=True):
            if u in keep and v in keep:
                g.add_edge(u, v, **d)
        return g

    def __getitem__(self, n):
        return self._adj[n]

    def __len__(self):
        return len(self._adj)

    def __iter__(self):
        return iter(self._adj)

    def __contains__(self, n):
        return n in self._adj


def degree_centrality(G):
    n = len(G)
    if n <= 1:
        return {v: 1.0 for v in G}
    return {v: len(G[v]) / (n - 1.0) for v in G}


def betweenness_centrality(G, normalized=True):
    bc = dict.fromkeys(G, 0.0)
    for s in G:
        stack, pred = [], {v: [] for v in G}
        sigma = dict.fromkeys(G, 0.0)
        dist = {}
        sigma[s] = 1.0
        dist[s] = 0
        q = deque([s])
        while q:
            v = q.popleft()
            stack.append(v)
            for w in G[v]:
                if w not in dist:
                    dist[w] = dist[v] + 1
                    q.append(w)
                if dist[w] == dist[v] + 1:
                    sigma[w] += sigma[v]
                    pred[w].append(v)
        delta = dict.fromkeys(G, 0.0)
        while stack:
            w = stack.pop()
            for v in pred[w]:
                delta[v] += sigma[v] / sigma[w] * (1.0 + delta[w])
            if w != s:
                bc[w] += delta[w]
    n = len(G)
    scale = (1.0 / ((n - 1) * (n - 2)) if n > 2 else None) if normalized else 0.5
    if scale is not None:
        for v in bc:
            bc[v] *= scale
    return bc


def spring_layout(G, **kw):
    return {v: (0.0, 0.0) for v in G}


def draw_networkx(*a, **kw):
    return None
