"""Picture sharding across the GPUs of one box: a picture's ME depends only on source pyramids, so
pictures are the unit. No collective on the search; ranks only need the pyramids of the pictures they reference."""


def pictures_for_rank(decode_order_pocs, rank, world):
    """Pictures of a (mini-GOP's) decode-order list that `rank` of `world` processes: dealt round-robin."""
    return list(decode_order_pocs[rank::world])


def pyramids_needed(pocs, ref_lists_of):
    """Every picture whose pyramid must be resident on a rank that processes `pocs`.
    ref_lists_of(poc) -> (list0 pocs, list1 pocs)"""
    need = set()
    for poc in pocs:
        l0, l1 = ref_lists_of(poc)
        need.update([poc], l0, l1)
    return sorted(need)


def owners(decode_order_pocs, world, cost=None, group=16):
    """picture -> rank that uploads it and runs its ME.
    cost None: the sequence's decode order dealt round-robin.
    cost(poc) -> number: within every mini-GOP (`group` pictures of the decode order, the unit the encoder's lookahead
    hands to ME at once) the most expensive picture first, each to the rank with the least work so far (longest
    processing time first; ties: lower rank): a base-layer picture with 2 references and a top-layer one with 7 differ by
    3.5x in search work, which round-robin dealing turns into idle GPUs. Deterministic: every rank computes the same map."""
    if cost is None:
        return {poc: i % world for i, poc in enumerate(decode_order_pocs)}
    own, load = {}, [0.0] * world
    order = list(decode_order_pocs)
    for g0 in range(0, len(order), group):
        chunk = sorted(order[g0:g0 + group], key=lambda poc: (-cost(poc), order.index(poc)))
        for poc in chunk:
            r = min(range(world), key=lambda k: (load[k], k))
            own[poc] = r
            load[r] += cost(poc)
    return own
