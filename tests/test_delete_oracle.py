"""Oracle restatement of Graph.Delete (graph.go:843-866) on hand-built layers: the expected lists below were derived by
hand from isolate (graph.go:223-235) and replenish (graph.go:172-219), quirks included."""
import numpy as np


def _layer(oracle, pts, links, M=2):
    g = oracle.Graph(2, M, oracle.EUCLIDEAN)
    for key, p in enumerate(pts):
        g.put_node(key, np.asarray(p, dtype=np.float32))
    for a, nbs in links.items():
        for b in nbs:
            g.link(0, a, b)
    return g


PTS = [(1, 0), (1, 0.1), (1, 0.2), (1, 0.5), (0, 1)]


def test_delete_replenishes_through_neighbours_of_neighbours(oracle):
    # A=0 {B, D}, B=1 {A, C}, C=2 {B, D}, D=3 {A, C}; delete B.
    g = _layer(oracle, PTS[:4], {0: [1, 3], 1: [0, 2], 2: [1, 3], 3: [0, 2]})
    assert g.delete(1) is True
    assert g.delete(1) is False          # second Delete of the same key
    assert g.delete(99) is False
    assert len(g) == 3 and g.is_dead(1)
    # A loses B, finds C through D (the only neighbour of a neighbour it does not know yet); C finds A the same way
    assert list(g.neighbors(0, 0)) == [2, 3]
    assert list(g.neighbors(0, 2)) == [0, 3]
    assert list(g.neighbors(0, 3)) == [0, 2]
    assert list(g.neighbors(0, 1)) == [0, 2]   # isolate never touches the deleted node's own map


def test_delete_can_leave_the_node_linked(oracle):
    # A=0 {B, C}, B=1 {A, C}, C=2 {A, B}: when A drops B, C still lists B, so replenish hands B straight back to A
    # (graph.go:193-210 only skips A itself and A's *current* neighbours); then the same happens to C through A.
    g = _layer(oracle, PTS[:3], {0: [1, 2], 1: [0, 2], 2: [0, 1]})
    assert g.delete(1) is True
    assert len(g) == 2 and g.is_dead(1)
    assert list(g.neighbors(0, 0)) == [1, 2]
    assert list(g.neighbors(0, 2)) == [0, 1]
    # ... and Search still returns the deleted node
    rows, d, cnt, _ = g.batch_search(np.asarray([PTS[1]], dtype=np.float32), 3, 20)
    assert cnt[0] == 3 and 1 in rows[0]


def test_delete_entry_point_moves_to_lowest_live_row(oracle):
    g = _layer(oracle, PTS, {0: [1], 1: [0, 2], 2: [1, 3], 3: [2, 4], 4: [3]})
    assert g.entries()[0] == 0
    assert g.delete(0)
    assert g.entries()[0] == 1
    assert g.batch_delete([1, 2, 3, 4]) == [True] * 4
    assert len(g) == 0 and g.entries()[0] == -1
    rows, d, cnt, _ = g.batch_search(np.asarray([PTS[0]], dtype=np.float32), 3, 20)
    assert cnt[0] == 0
