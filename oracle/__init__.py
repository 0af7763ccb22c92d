"""CPU oracle for the vllm_router routing hot path.

TEST INFRASTRUCTURE ONLY.  The product package (vllm-production-stack_b200/)
never imports this; only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs do, as the checker.

Modules
-------
coracle     ctypes front-end of oracle.c (fast: xxh64, pointer trie, md5, ring)
prefix_ref  Python restatement of HashTrie + PrefixAwareRouter.route_request,
            asyncio locks included, used as the timed "reference arm"
ring_ref    Python restatement of uhashring==2.3's default ring + SessionRouter
chooser     the canonical tie-break shared by both sides

Parity pinning: the reference ships no golden vectors for this path
(SURVEY.md 8c).  The oracle is pinned by tests/golden/*.json, generated in the
build container by scripts/gen_golden.py from the UNMODIFIED reference
(vllm_router.prefix.hashtrie.HashTrie and PrefixAwareRouter.route_request) and
by the XXH64 / MD5 known answers.  The consistent-hash ring stays "parity
unpinned": uhashring is absent from the reference tree and not installable
here, so ring_ref is a restatement of its published algorithm only.
"""
