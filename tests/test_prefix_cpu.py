"""Prefix KV reuse, host logic, property-tested without a GPU through dis_prefix_index_*.

The reference specifies (.kiro/specs/distributed-inference-server/design.md:734-744)
  Property 9  -- a request whose tokens start with a cached request's tokens reuses the cached K/V for the shared
                 prefix;
  Property 10 -- under memory pressure the least recently used entries are evicted first.
The model below restates those rules for a PAGED cache in plain Python (a registry keyed by the token prefix, an
LRU list of pages nobody maps) and every operation on the C++ index is checked against it: which pages are reused,
which are evicted and in which order, when admission is refused, and the CacheStats counters."""
import ctypes as C

import numpy as np
from hypothesis import given, settings, strategies as st

from distributed_inference_server_b200 import binding as B


class Model:
    def __init__(self, ps, n_pages):
        self.ps, self.free = ps, set(range(n_pages))
        self.registry = {}      # token prefix (tuple, whole pages) -> page
        self.key_of = {}        # page -> its prefix
        self.refs = {}
        self.lru = []           # front = next victim
        self.evictions = self.hits = self.hit_tokens = self.lookups = self.lookup_tokens = 0

    def release(self, pages):
        for p in reversed(pages):
            if p in self.key_of:
                self.refs[p] -= 1
                if self.refs[p] == 0:
                    self.lru.append(p)
            else:
                self.free.add(p)

    def admit(self, tokens):
        """-> None if refused, else (matched pages, number of pages from the free set, victims in order)"""
        n, ps = len(tokens), self.ps
        need = (n + 1 + ps - 1) // ps
        self.lookups += 1
        self.lookup_tokens += n
        matched = []
        for i in range((n - 1) // ps):
            p = self.registry.get(tuple(tokens[:(i + 1) * ps]))
            if p is None:
                break
            if self.refs[p] == 0:
                self.lru.remove(p)
            self.refs[p] += 1
            matched.append(p)
        if matched:
            self.hits += 1
            self.hit_tokens += len(matched) * ps
        k_new = need - len(matched)
        if k_new > len(self.free) + len(self.lru):
            self.release(matched)
            return None
        from_free = min(k_new, len(self.free))
        victims = self.lru[:k_new - from_free]
        for v in victims:
            self.lru.remove(v)
            del self.registry[self.key_of.pop(v)]
            del self.refs[v]
            self.evictions += 1
        return matched, from_free, victims

    def commit(self, tokens, n_tokens, pages):
        for i in range(min(n_tokens // self.ps, len(pages))):
            key, p = tuple(tokens[:(i + 1) * self.ps]), pages[i]
            if p in self.key_of or key in self.registry:
                continue
            self.registry[key], self.key_of[p], self.refs[p] = p, key, 1


def _arr(v, dt):
    return np.ascontiguousarray(v, dt)


ops = st.lists(st.tuples(st.sampled_from(["admit", "admit", "commit", "release"]),
                         st.lists(st.integers(0, 1), min_size=1, max_size=14), st.integers(0, 1 << 30)), min_size=1, max_size=60)


@settings(max_examples=300, deadline=None)
@given(ps=st.sampled_from([1, 2, 4]), n_pages=st.integers(3, 12), script=ops)
def test_index_follows_the_paged_restatement_of_properties_9_and_10(ps, n_pages, script):
    lib = B.load()
    h = C.c_void_p()
    assert lib.dis_prefix_index_create(ps, n_pages, C.byref(h)) == B.OK
    model, live = Model(ps, n_pages), []   # live: [tokens, pages]
    try:
        for op, tokens, pick in script:
            if op == "admit":
                t = _arr(tokens, np.uint32)
                pages = np.zeros(32, np.int32)
                n_out, matched = C.c_int32(), C.c_int32()
                rc = lib.dis_prefix_index_admit(h, t.ctypes.data, len(tokens), pages.ctypes.data, 32, C.byref(n_out), C.byref(matched))
                want = model.admit(tokens)
                if want is None:
                    assert rc == B.ERR_OUT_OF_MEMORY
                else:
                    assert rc == B.OK
                    m_pages, from_free, victims = want
                    got = [int(p) for p in pages[:n_out.value]]
                    assert n_out.value == (len(tokens) + 1 + ps - 1) // ps
                    assert matched.value == len(m_pages) * ps                      # Property 9: the shared prefix ...
                    assert got[:len(m_pages)] == m_pages                           # ... maps the very same pages
                    fresh = got[len(m_pages):]
                    assert set(fresh[:from_free]) <= model.free                    # free pages first,
                    model.free -= set(fresh[:from_free])
                    assert fresh[from_free:] == victims                            # then Property 10: LRU order
                    held = [p for _, ps_ in live for p in ps_]
                    assert not (set(fresh) & set(held)) and len(set(got)) == len(got)
                    live.append([tokens, got])
            elif op == "commit" and live:
                tokens, pages = live[pick % len(live)]
                n_tok = pick % (len(tokens) + 1)
                assert lib.dis_prefix_index_commit(h, _arr(tokens, np.uint32).ctypes.data, n_tok,
                                                   _arr(pages, np.int32).ctypes.data, len(pages)) == B.OK
                model.commit(tokens, n_tok, pages)
            elif op == "release" and live:
                tokens, pages = live.pop(pick % len(live))
                assert lib.dis_prefix_index_release(h, _arr(pages, np.int32).ctypes.data, len(pages)) == B.OK
                model.release(pages)
            s = B.PrefixStats()
            assert lib.dis_prefix_index_stats(h, C.byref(s)) == B.OK
            assert (s.cached_pages, s.evictable_pages, s.free_pages) == (len(model.registry), len(model.lru), len(model.free))
            assert (s.lookups, s.hits, s.lookup_tokens, s.hit_tokens, s.evictions) == \
                   (model.lookups, model.hits, model.lookup_tokens, model.hit_tokens, model.evictions)
            distinct_live = {p for _, ps_ in live for p in ps_}
            assert s.free_pages + s.evictable_pages + len(distinct_live) == n_pages   # no page is lost or duplicated
    finally:
        lib.dis_prefix_index_destroy(h)


def test_reuse_leaves_at_least_one_token_to_compute_and_deeper_pages_go_first():
    lib = B.load()
    h = C.c_void_p()
    assert lib.dis_prefix_index_create(4, 6, C.byref(h)) == B.OK
    toks = _arr(list(range(12)), np.uint32)       # exactly 3 pages
    pages = np.zeros(8, np.int32)
    n, m = C.c_int32(), C.c_int32()
    assert lib.dis_prefix_index_admit(h, toks.ctypes.data, 12, pages.ctypes.data, 8, C.byref(n), C.byref(m)) == B.OK
    first = [int(p) for p in pages[:n.value]]
    assert (n.value, m.value) == (4, 0)
    assert lib.dis_prefix_index_commit(h, toks.ctypes.data, 12, pages.ctypes.data, n.value) == B.OK
    assert lib.dis_prefix_index_release(h, pages.ctypes.data, n.value) == B.OK
    # the same 12 tokens again: the last page is recomputed so that the request still produces logits
    assert lib.dis_prefix_index_admit(h, toks.ctypes.data, 12, pages.ctypes.data, 8, C.byref(n), C.byref(m)) == B.OK
    assert m.value == 8 and [int(p) for p in pages[:2]] == first[:2]
    assert lib.dis_prefix_index_release(h, pages.ctypes.data, n.value) == B.OK
    # a stranger needing 5 pages: 3 free + the two deepest cached pages (page 2 of the chain, then page 1)
    other = _arr(list(range(100, 119)), np.uint32)
    assert lib.dis_prefix_index_admit(h, other.ctypes.data, 19, pages.ctypes.data, 8, C.byref(n), C.byref(m)) == B.OK
    got = [int(p) for p in pages[:n.value]]
    assert m.value == 0 and first[2] in got and first[1] in got and first[0] not in got
    s = B.PrefixStats()
    lib.dis_prefix_index_stats(h, C.byref(s))
    assert (s.evictions, s.cached_pages, s.evictable_pages) == (2, 1, 1)   # what is left is still a usable prefix
    lib.dis_prefix_index_destroy(h)
