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


def owners(decode_order_pocs, world):
    """picture -> rank that uploads it and runs its ME: the sequence's decode order dealt round-robin"""
    return {poc: i % world for i, poc in enumerate(decode_order_pocs)}
