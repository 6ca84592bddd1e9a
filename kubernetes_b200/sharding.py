"""Row ranges of a sharded node table (mirror of Engine::plan_shard in csrc/engine/engine_run.cpp) and the
handle exchange that connects one engine per GPU. The scheduling stream itself is replicated: every rank
feeds every node and pod to its engine; only device memory is partitioned."""


def rows_per_shard(n_nodes, n_ranks):
    if n_ranks <= 1:
        return max(256, (n_nodes + 255) & ~255)
    per = (n_nodes + n_ranks - 1) // n_ranks
    return max(256, (per + 255) & ~255)  # whole 256-row segments per shard


def plan(n_nodes, n_ranks):
    """[(first_row, row_count)] per rank; contiguous in bytewise node-name order, so the round-robin tie order of
    selectHost (descending name) is descending (rank, row)."""
    rps = rows_per_shard(n_nodes, n_ranks)
    out = []
    for r in range(n_ranks):
        base = min(n_nodes, r * rps)
        out.append((base, max(0, min(n_nodes - base, rps))))
    return out


def connect(engine, rank, world, all_gather_object, max_batch_pods):
    """Export this rank's CUDA-IPC handles, gather everybody's, connect. `all_gather_object(list, obj)` is
    torch.distributed.all_gather_object (any backend)."""
    blob = engine.shard_export(max_batch_pods)
    blobs = [None] * world
    all_gather_object(blobs, blob)
    engine.shard_connect(blobs)
    return blobs
