"""Table -> rank placement policies, same registry surface as the reference (sharders.py:1-62):
`shard(Es, ndevices, alg)` returns the device index of every table; `register_sharder(name)` adds one.

Host-side integer logic on the hot path's boundary: results must equal the
reference's bit for bit (tests/test_host_cpu.py checks them against tests/golden/sharders.npz, which
the reference itself produced).
"""
_sharders = {}


def get_splits(T, ndevices):
    """Contiguous chunk sizes: the first T % ndevices ranks get one extra table (sharders.py:3-9)."""
    base, extra = divmod(T, ndevices)
    return [base + (1 if r < extra else 0) for r in range(ndevices)]


def register_sharder(sharder_name):
    def decorate(func):
        _sharders[sharder_name] = func
        return func
    return decorate


def shard(Es, ndevices, alg="naive"):
    if alg not in _sharders:
        raise ValueError("ERROR: sharder not found: %r (have %s)" % (alg, sorted(_sharders)))
    return _sharders[alg](Es, ndevices)


@register_sharder("naive")
def round_robin_shard(Es, ndevices):
    return [t % ndevices for t in range(len(Es))]


@register_sharder("naive_chunk")
def chunk_shard(Es, ndevices):
    out = []
    for rank, count in enumerate(get_splits(len(Es), ndevices)):
        out += [rank] * count
    return out


@register_sharder("greedy")
def greedy_shard(Es, ndevices):
    """Tables in order; each goes to the rank holding the fewest rows so far, ties to the lowest rank."""
    rows_on = [0] * ndevices
    out = []
    for E in Es:
        rank = min(range(ndevices), key=lambda r: (rows_on[r], r))
        rows_on[rank] += int(E)
        out.append(rank)
    return out


@register_sharder("hardcode")
def hardcode_shard(Es, ndevices):
    return [0] + [1] * (len(Es) - 1)


@register_sharder("traffic")
def traffic_shard(Es, ndevices):
    """Not in the reference: balances the NUMBER of tables per rank first (the
    all-to-all egress and lookup traffic are per table, not per row) and rows second: tables sorted by
    rows descending are dealt to the rank with the fewest tables, ties to the fewest rows."""
    order = sorted(range(len(Es)), key=lambda t: (-int(Es[t]), t))
    count = [0] * ndevices
    rows_on = [0] * ndevices
    out = [0] * len(Es)
    for t in order:
        rank = min(range(ndevices), key=lambda r: (count[r], rows_on[r], r))
        count[rank] += 1
        rows_on[rank] += int(Es[t])
        out[t] = rank
    return out
