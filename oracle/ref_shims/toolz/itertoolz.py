import itertools

_no_default = "__no__default__"
_no_pad = "__no__pad__"


def unique(seq, key=None):
    seen = set()
    for item in seq:
        val = item if key is None else key(item)
        if val not in seen:
            seen.add(val)
            yield item


def groupby(key, seq):
    if not callable(key):
        k = key
        key = (lambda x: tuple(x[i] for i in k)) if isinstance(k, list) else (lambda x: x[k])
    out = {}
    for item in seq:
        out.setdefault(key(item), []).append(item)
    return out


def reduceby(key, binop, seq, init=_no_default):
    if not callable(key):
        k = key
        key = (lambda x: tuple(x[i] for i in k)) if isinstance(k, list) else (lambda x: x[k])
    out = {}
    for item in seq:
        kk = key(item)
        if kk not in out:
            if init is _no_default:
                out[kk] = item
                continue
            out[kk] = init() if callable(init) else init
        out[kk] = binop(out[kk], item)
    return out


def pluck(ind, seqs, default=_no_default):
    if isinstance(ind, list):
        if default is _no_default:
            return (tuple(s[i] for i in ind) for s in seqs)
        return (tuple(_get(s, i, default) for i in ind) for s in seqs)
    if default is _no_default:
        return (s[ind] for s in seqs)
    return (_get(s, ind, default) for s in seqs)


def _get(s, i, default):
    try:
        return s[i]
    except (KeyError, IndexError, TypeError):
        return default


def take_nth(n, seq):
    return itertools.islice(seq, 0, None, n)


def partition(n, seq, pad=_no_pad):
    args = [iter(seq)] * n
    if pad is _no_pad:
        return zip(*args)
    return itertools.zip_longest(*args, fillvalue=pad)


def partition_all(n, seq):
    it = iter(seq)
    while True:
        chunk = tuple(itertools.islice(it, n))
        if not chunk:
            return
        yield chunk


def diff(*seqs, **kwargs):
    default = kwargs.get("default", _no_default)
    key = kwargs.get("key")
    if len(seqs) == 1 and isinstance(seqs[0], list):
        seqs = seqs[0]
    if default is _no_default:
        iters = zip(*seqs)
    else:
        iters = itertools.zip_longest(*seqs, fillvalue=default)
    for items in iters:
        vals = items if key is None else tuple(map(key, items))
        if vals.count(vals[0]) != len(vals):
            yield items
