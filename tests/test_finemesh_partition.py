"""Host logic of the partitioned fine-mesh trajectory (finemesh_dist.node_partition): CPU only."""
import pytest


def test_node_partition_blocks_cover_the_mesh_in_whole_cta_blocks():
    from burgers_b200.finemesh_dist import node_partition, BLOCK_NODES
    for N, R in [(1 << 20, 1), (1 << 20, 2), (1 << 20, 8), (100003, 2), (100003, 3), (8192 * 5, 5), (131079, 8),
                 (8 * 148 * 8192, 8), (8192 * 139 - 4321, 2)]:
        parts = node_partition(N, R)
        assert len(parts) == R and parts[0][0] == 0
        assert sum(c for _, c in parts) == N
        for r, (off, cnt) in enumerate(parts):
            assert cnt >= 1 and off % BLOCK_NODES == 0
            if r < R - 1:
                assert cnt % BLOCK_NODES == 0 and parts[r + 1][0] == off + cnt
        ctas = [-(-c // BLOCK_NODES) for _, c in parts]
        assert max(ctas) - min(ctas) <= 1


def test_node_partition_rejects_empty_ranks_and_too_many_ranks():
    from burgers_b200.finemesh_dist import node_partition
    with pytest.raises(ValueError):
        node_partition(8192 * 3, 4)
    with pytest.raises(ValueError):
        node_partition(1 << 20, 9)
    with pytest.raises(ValueError):
        node_partition(1 << 20, 0)
