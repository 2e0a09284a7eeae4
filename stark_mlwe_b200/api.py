"""Host-side mirror of the reference's Rust API for the commitment hot path, over the C ABI of libsmw.so.

Names and argument meaning follow the reference crates (file:line cited per item) so the parity tests read
like the reference's own tests; conditions on which the reference panics raise SmwError here (the C ABI
returns a status code, include/smw.h).  Field elements travel as numpy uint64 arrays of shape (..., 4)
(Montgomery limbs, the memory image of ark-ff's Fp256) or as torch CUDA tensors of dtype int64 with the
same layout, which are used in place on the device.  All arithmetic happens in the CUDA library.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import HOST, DEVICE

P = 0x40000000000000000000000000000000224698FC0994A8DD8C46EB2100000001


class SmwError(RuntimeError):
    def __init__(self, code, msg=""):
        self.code = code
        self.name = _lib.ERR.get(code, str(code))
        super().__init__("%s: %s" % (self.name, msg))


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _buf(x):
    """-> (pointer, n_fr, mem, keepalive)"""
    if x is None:
        return None, 0, HOST, None
    if _is_torch(x):
        assert x.is_cuda and x.is_contiguous() and x.element_size() == 8 and x.shape[-1] == 4
        return C.c_void_p(x.data_ptr()), x.numel() // 4, DEVICE, x
    a = np.ascontiguousarray(x, dtype=np.uint64)
    assert a.shape[-1] == 4, "field arrays have shape (..., 4)"
    return C.c_void_p(a.ctypes.data), a.size // 4, HOST, a


def _fr1(x):
    a = np.ascontiguousarray(x, dtype=np.uint64).reshape(4)
    return a


def _out(n):
    return np.zeros((n, 4), dtype=np.uint64)


def _bytes(b):
    b = bytes(b)
    return (C.c_uint8 * max(len(b), 1)).from_buffer_copy(b if b else b"\0"), len(b)


class Context:
    """One CUDA device + stream (smw_ctx). Not in the reference: the reference has no device."""

    def __init__(self, device=0, stream=None):
        self.L = _lib.load()
        h = C.c_void_p()
        # stream: None = the library creates its own (non-blocking) stream; a cudaStream_t handle = run there.  The handle 0 is the legacy
        # default stream (what torch.cuda.current_stream().cuda_stream returns until a stream is set): it is passed as cudaStreamLegacy
        # (0x1), NOT as "none" — a private non-blocking stream does not order with work the caller enqueues on the default stream.
        sp = None if stream is None else C.c_void_p(int(stream) if int(stream) != 0 else 1)
        rc = self.L.smw_ctx_create_on_stream(int(device), sp, C.byref(h))
        if rc:
            raise SmwError(rc, "cannot create a context on CUDA device %d (no CPU fallback exists)" % device)
        self.h = h
        self.device = device

    def check(self, rc):
        if rc:
            raise SmwError(rc, (self.L.smw_last_error(self.h) or b"").decode())

    def sync(self):
        self.check(self.L.smw_ctx_sync(self.h))

    def release_cached(self):
        """give the idle blocks of the context's arena pool (large temporaries kept between calls) back to the driver"""
        self.check(self.L.smw_ctx_release_cached(self.h))

    def launch_count(self):
        return int(self.L.smw_ctx_launch_count(self.h))

    def stream(self):
        return self.L.smw_ctx_stream(self.h)

    def collective_count(self):
        return int(self.L.smw_ctx_collective_count(self.h))

    def tc_node_counts(self, params):
        """{fp64_instr, int8_mac, word_products, launches_per_batch, group, min_nodes} of one arity-16 node on the tensor-core path"""
        out = (C.c_uint64 * 6)()
        self.check(self.L.smw_diag_tc_node_counts(self.h, params.h, out))
        return dict(zip(("fp64_instr", "int8_mac", "word_products", "launches_per_batch", "group", "min_nodes"), [int(v) for v in out]))

    def set_tc_min_nodes(self, n):
        """Merkle levels of >= n full arity-16 nodes run their linear layers on the INT8 tensor cores (0 = never)"""
        self.check(self.L.smw_ctx_set_tc_min_nodes(self.h, int(n)))

    KERNEL_CLASSES = ("merkle_level", "pair_leaves", "fri_leaf", "permute", "ntt_pass", "lde_aux", "fold", "misc")

    def profile(self, on=True):
        """per-kernel-class device timing (CUDA events on the context's stream); a measurement aid, not in the reference"""
        self.check(self.L.smw_ctx_profile_enable(self.h, 1 if on else 0))

    def profile_read(self):
        """-> {class: (ms, launches)} since the last read; synchronises the stream"""
        n = len(self.KERNEL_CLASSES)
        ms, cnt = (C.c_double * n)(), (C.c_uint64 * n)()
        self.check(self.L.smw_ctx_profile_read(self.h, ms, cnt, n))
        return {k: (ms[i], int(cnt[i])) for i, k in enumerate(self.KERNEL_CLASSES)}

    def mac_rate(self):
        """measured IMAD.WIDE.U32 rate: (32x32+64 multiply-accumulates per second, warp instructions per clock per SM)"""
        a, b = C.c_double(), C.c_double()
        self.check(self.L.smw_diag_mac_rate(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def fp64_rate(self):
        """measured DFMA.RZ rate: (instructions per second over all lanes, warp instructions per clock per SM)"""
        a, b = C.c_double(), C.c_double()
        self.check(self.L.smw_diag_fp64_rate(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def pipe_rate(self, fp64):
        """live pipe-rate measurement with its own clock record: lanes_per_s = warp_instr_per_clk_per_sm x 32 x n_sm x sm_hz_during"""
        o = (C.c_double * 5)()
        self.check(self.L.smw_diag_pipe_rate(self.h, 1 if fp64 else 0, o))
        return {"pipe": "fp64 DFMA.RZ" if fp64 else "IMAD.WIDE.U32 (32x32+64)", "lanes_per_s": o[0], "warp_instr_per_clk_per_sm": o[1], "sm_hz_during": o[2],
                "n_sm": int(o[3]), "sm_max_hz": o[4], "check_lanes_per_s": o[1] * 32 * o[3] * o[2]}

    def close(self):
        if self.h:
            self.L.smw_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- field helpers (ark-ff / field/src/lib.rs:206-215)
    def from_canonical(self, canon):
        p, n, mem, _k = _buf(canon)
        out = _out(n)
        self.check(self.L.smw_fr_from_canonical(self.h, p, C.c_void_p(out.ctypes.data), n, HOST))
        return out

    def to_canonical(self, fr):
        p, n, mem, _k = _buf(fr)
        out = _out(n)
        self.check(self.L.smw_fr_to_canonical(self.h, p, C.c_void_p(out.ctypes.data), n, HOST))
        return out

    def root_of_unity(self, log_n):
        out = _out(1)
        self.check(self.L.smw_root_of_unity(self.h, int(log_n), C.c_void_p(out.ctypes.data)))
        return out[0]


# ----------------------------------------------------------------------------- poseidon
class PoseidonParamsDynamic:
    """poseidon/src/lib.rs:104-114 (device resident; static t=17 `PoseidonParams` is the same object with t == 17)."""

    def __init__(self, ctx, h, owned):
        self.ctx, self.h, self.owned = ctx, h, owned
        self.t = int(ctx.L.smw_params_width(h))
        self.rate = self.t - 1
        self.rounds_full = 8
        self.rounds_partial = int(ctx.L.smw_params_partial_rounds(h))

    def executed_macs(self):
        """32x32+64 multiply-accumulates one production permutation executes (counted from the device program)"""
        m = C.c_uint64()
        self.ctx.check(self.ctx.L.smw_diag_perm_macs(self.h, C.byref(m)))
        return int(m.value)

    def executed_fp64_ops(self):
        """FP64-pipe instructions (DFMA + DADD) of one production permutation on the fr52 multiplier; 0 on the integer one"""
        m = C.c_uint64()
        self.ctx.check(self.ctx.L.smw_diag_perm_fp64_ops(self.h, C.byref(m)))
        return int(m.value)

    def export(self):
        t, rp = self.t, self.rounds_partial
        mds, rcf, rcp = np.zeros((t, t, 4), np.uint64), np.zeros((8, t, 4), np.uint64), np.zeros((rp, 4), np.uint64)
        self.ctx.check(self.ctx.L.smw_params_export(self.h, C.c_void_p(mds.ctypes.data), C.c_void_p(rcf.ctypes.data), C.c_void_p(rcp.ctypes.data)))
        return mds, rcf, rcp

    def __del__(self):
        try:
            if self.owned and self.ctx.h:
                self.ctx.L.smw_params_free(self.h)
        except Exception:
            pass


def generate_params_t17_x5(ctx, seed):
    """poseidon::params::generate_params_t17_x5 (poseidon/src/lib.rs:318-356)"""
    buf, n = _bytes(seed)
    h = C.c_void_p()
    ctx.check(ctx.L.smw_params_from_seed(ctx.h, buf, n, 17, 64, C.byref(h)))
    return PoseidonParamsDynamic(ctx, h, True)


def params_from_seed(ctx, seed, t, rp):
    buf, n = _bytes(seed)
    h = C.c_void_p()
    ctx.check(ctx.L.smw_params_from_seed(ctx.h, buf, n, int(t), int(rp), C.byref(h)))
    return PoseidonParamsDynamic(ctx, h, True)


def params_from_tables(ctx, mds, rc_full, rc_partial):
    """A caller-built PoseidonParamsDynamic (MerkleChannelCfg::with_params accepts any, merkle/src/lib.rs:91)."""
    t, rp = np.asarray(mds).shape[0], np.asarray(rc_partial).reshape(-1, 4).shape[0]
    a, b, c = (np.ascontiguousarray(x, dtype=np.uint64) for x in (mds, rc_full, rc_partial))
    h = C.c_void_p()
    ctx.check(ctx.L.smw_params_from_tables(ctx.h, t, rp, C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data), C.c_void_p(c.ctypes.data), HOST, C.byref(h)))
    return PoseidonParamsDynamic(ctx, h, True)


def poseidon_params_for_width(ctx, t):
    """poseidon/src/lib.rs:120-146; panics (SMW_ERR_UNSUPPORTED_WIDTH) outside {9,17,33,65,129}"""
    h = C.c_void_p()
    ctx.check(ctx.L.smw_params_for_width(ctx.h, int(t), C.byref(h)))
    return PoseidonParamsDynamic(ctx, h, False)


def poseidon_params_for_arity(ctx, arity):
    """poseidon/src/lib.rs:155-166"""
    h = C.c_void_p()
    ctx.check(ctx.L.smw_params_for_arity(ctx.h, int(arity), C.byref(h)))
    return PoseidonParamsDynamic(ctx, h, False)


def commitment_default_params(ctx):
    """commitment::default_params (commitment/src/lib.rs:48-51) == merkle::default_params (merkle/src/lib.rs:365-368)"""
    h = C.c_void_p()
    ctx.check(ctx.L.smw_params_commitment_default(ctx.h, C.byref(h)))
    return PoseidonParamsDynamic(ctx, h, False)


def transcript_default_params(ctx):
    """transcript::default_params (transcript/src/lib.rs:44-46)"""
    h = C.c_void_p()
    ctx.check(ctx.L.smw_params_transcript_default(ctx.h, C.byref(h)))
    return PoseidonParamsDynamic(ctx, h, False)


def permute(ctx, state, params, dense=False):
    """permute / permute_dynamic (poseidon/src/lib.rs:31-68, 219-258). state: (batch, t, 4) or (t, 4); returns a permuted copy
    (torch CUDA tensors are permuted in place)."""
    if _is_torch(state):
        p, n, mem, _k = _buf(state)
        fn = ctx.L.smw_poseidon_permute_dense if dense else ctx.L.smw_poseidon_permute
        ctx.check(fn(ctx.h, params.h, p, n // params.t, DEVICE))
        return state
    s = np.ascontiguousarray(state, dtype=np.uint64).copy()
    assert s.shape[-2] == params.t, "state width must equal params.t"
    fn = ctx.L.smw_poseidon_permute_dense if dense else ctx.L.smw_poseidon_permute
    ctx.check(fn(ctx.h, params.h, C.c_void_p(s.ctypes.data), s.size // (4 * params.t), HOST))
    return s


def hash_with_ds_dynamic(ctx, ds_fields, inputs, params):
    """poseidon/src/lib.rs:288-312"""
    ds = np.ascontiguousarray(ds_fields, dtype=np.uint64).reshape(-1, 4)
    inp = np.ascontiguousarray(inputs, dtype=np.uint64).reshape(-1, 4)
    out = _out(1)
    ctx.check(ctx.L.smw_hash_with_ds_dynamic(ctx.h, params.h, C.c_void_p(ds.ctypes.data), ds.shape[0], C.c_void_p(inp.ctypes.data) if inp.size else None, inp.shape[0], C.c_void_p(out.ctypes.data), HOST))
    return out[0]


def hash_with_ds(ctx, inputs, ds_tag, params):
    """legacy sponge, poseidon/src/lib.rs:85-100"""
    inp = np.ascontiguousarray(inputs, dtype=np.uint64).reshape(-1, 4)
    tag = _fr1(ds_tag)
    out = _out(1)
    ctx.check(ctx.L.smw_hash_with_ds(ctx.h, params.h, C.c_void_p(inp.ctypes.data) if inp.size else None, inp.shape[0], C.c_void_p(tag.ctypes.data), C.c_void_p(out.ctypes.data), HOST))
    return out[0]


# ----------------------------------------------------------------------------- merkle
class MerkleChannelCfg:
    """merkle/src/lib.rs:84-112"""

    def __init__(self, ctx, arity, params, tree_label=0):
        self.ctx, self.arity, self.params, self.tree_label = ctx, arity, params, tree_label

    @classmethod
    def new(cls, ctx, arity):
        return cls(ctx, arity, poseidon_params_for_arity(ctx, arity), 0)

    @classmethod
    def with_params(cls, ctx, arity, params):
        return cls(ctx, arity, params, 0)

    def with_tree_label(self, label):
        return MerkleChannelCfg(self.ctx, self.arity, self.params, int(label))


class MerkleProof:
    """merkle/src/lib.rs:131-143"""

    @classmethod
    def from_fields(cls, indices, siblings, group_sizes, arity):
        """a proof as it arrives at a verifier (deserialised fields, no device handle)"""
        pr = cls.__new__(cls)
        pr.indices = [int(i) for i in indices]
        pr.siblings = [np.ascontiguousarray(x, dtype=np.uint64).reshape(-1, 4) for x in siblings]
        pr.group_sizes = [np.ascontiguousarray(g, dtype=np.uint8).reshape(-1) for g in group_sizes]
        pr.arity, pr.nlev = int(arity), len(pr.siblings)
        return pr

    def __init__(self, ctx, h, arity):
        L = ctx.L
        nl = int(L.smw_proof_num_levels(h))
        n = int(L.smw_proof_num_indices(h))
        idx = np.zeros(max(n, 1), dtype=np.uint64)
        L.smw_proof_indices(h, C.c_void_p(idx.ctypes.data))
        self.indices = [int(v) for v in idx[:n]]
        self.siblings, self.group_sizes = [], []
        for l in range(nl):
            ns, ng = int(L.smw_proof_level_siblings(h, l)), int(L.smw_proof_level_groups(h, l))
            sib, grp = np.zeros((max(ns, 1), 4), np.uint64), np.zeros(max(ng, 1), np.uint8)
            L.smw_proof_level(h, l, C.c_void_p(sib.ctypes.data), C.c_void_p(grp.ctypes.data))
            self.siblings.append(sib[:ns].copy())
            self.group_sizes.append(grp[:ng].copy())
        self.arity = arity
        self.nlev = nl
        L.smw_proof_free(h)


def _verify_call(fn, root, indices, payload, per, proof, tree_label, dyn_params):
    ctx = dyn_params.ctx
    idx = np.ascontiguousarray(indices, dtype=np.uint64).reshape(-1)
    pay = np.ascontiguousarray(payload, dtype=np.uint64).reshape(-1, per * 4)
    if idx.size != pay.shape[0]:
        return False                                    # indices.len() != values.len()
    if len(proof.siblings) != len(proof.group_sizes):
        return False                                    # merkle/src/lib.rs:604
    nl = len(proof.siblings)
    pidx = np.ascontiguousarray(proof.indices, dtype=np.uint64).reshape(-1)
    nsib = (C.c_size_t * max(nl, 1))(*[int(x.shape[0]) for x in proof.siblings])
    ngrp = (C.c_size_t * max(nl, 1))(*[int(g.shape[0]) for g in proof.group_sizes])
    sib = np.ascontiguousarray(np.concatenate([x.reshape(-1, 4) for x in proof.siblings] + [np.zeros((1, 4), np.uint64)]))
    grp = np.ascontiguousarray(np.concatenate([g for g in proof.group_sizes] + [np.zeros(1, np.uint8)]))
    rt = _fr1(root)
    ok = C.c_int32(0)
    ctx.check(fn(ctx.h, dyn_params.h, int(proof.arity), int(tree_label), C.c_void_p(rt.ctypes.data),
                 C.c_void_p(idx.ctypes.data) if idx.size else None, C.c_void_p(pay.ctypes.data) if idx.size else None, idx.size,
                 C.c_void_p(pidx.ctypes.data) if pidx.size else None, pidx.size, nl, nsib, C.c_void_p(sib.ctypes.data), ngrp, C.c_void_p(grp.ctypes.data), C.byref(ok)))
    return bool(ok.value)


def verify_many_ds(root, indices, values, proof, tree_label, dyn_params):
    """merkle/src/lib.rs:587-701 — same argument order and the same `false` cases; parents are recomputed on the device"""
    return _verify_call(dyn_params.ctx.L.smw_verify_many_ds, root, indices, values, 1, proof, tree_label, dyn_params)


def verify_pairs_ds(root, indices, pairs, proof, tree_label, dyn_params):
    """merkle/src/lib.rs:725-773: `pairs` is n x (f, cp); leaf digests use LEAF_LEVEL_DS"""
    return _verify_call(dyn_params.ctx.L.smw_verify_pairs_ds, root, indices, pairs, 2, proof, tree_label, dyn_params)


def _verify_legacy_call(ctx, fn, root, indices, payload, per, proof, ds_tag, params):
    idx = np.ascontiguousarray(indices, dtype=np.uint64).reshape(-1)
    pay = np.ascontiguousarray(payload, dtype=np.uint64).reshape(-1, per * 4)
    if idx.size != pay.shape[0] or len(proof.siblings) != len(proof.group_sizes):
        return False                                    # merkle/src/lib.rs:495, 505
    nl = len(proof.siblings)
    pidx = np.ascontiguousarray(proof.indices, dtype=np.uint64).reshape(-1)
    nsib = (C.c_size_t * max(nl, 1))(*[int(x.shape[0]) for x in proof.siblings])
    ngrp = (C.c_size_t * max(nl, 1))(*[int(g.shape[0]) for g in proof.group_sizes])
    sib = np.ascontiguousarray(np.concatenate([x.reshape(-1, 4) for x in proof.siblings] + [np.zeros((1, 4), np.uint64)]))
    grp = np.ascontiguousarray(np.concatenate([g for g in proof.group_sizes] + [np.zeros(1, np.uint8)]))
    rt, tag = _fr1(root), _fr1(ds_tag)
    ok = C.c_int32(0)
    ctx.check(fn(ctx.h, params.h if params else None, int(proof.arity), C.c_void_p(tag.ctypes.data), C.c_void_p(rt.ctypes.data),
                 C.c_void_p(idx.ctypes.data) if idx.size else None, C.c_void_p(pay.ctypes.data) if idx.size else None, idx.size,
                 C.c_void_p(pidx.ctypes.data) if pidx.size else None, pidx.size, nl, nsib, C.c_void_p(sib.ctypes.data), ngrp, C.c_void_p(grp.ctypes.data), C.byref(ok)))
    return bool(ok.value)


def verify_many(ctx, root, indices, values, proof, ds_tag, params=None):
    """merkle/src/lib.rs:487-584 (legacy: parents are hash_with_ds(children, ds_tag))"""
    return _verify_legacy_call(ctx, ctx.L.smw_verify_many_legacy, root, indices, values, 1, proof, ds_tag, params)


def verify_pairs_legacy(ctx, root, indices, pairs, proof, ds_tag, params=None):
    """merkle/src/lib.rs:704-721"""
    return _verify_legacy_call(ctx, ctx.L.smw_verify_pairs_legacy, root, indices, pairs, 2, proof, ds_tag, params)


class MerkleTree:
    """merkle/src/lib.rs:114-128, 147-193, 392-445. `levels[0]` are the leaves."""

    def __init__(self, ctx, h, cfg, owned=True):
        self.ctx, self.h, self.cfg, self.owned = ctx, h, cfg, owned

    @classmethod
    def new(cls, leaves, cfg):
        ctx = cfg.ctx
        p, n, mem, _k = _buf(leaves)
        h = C.c_void_p()
        ctx.check(ctx.L.smw_merkle_new(ctx.h, cfg.params.h, int(cfg.arity), int(cfg.tree_label), p, n, mem, C.byref(h)))
        return cls(ctx, h, cfg)

    @classmethod
    def new_pairs(cls, f_vals, cp_vals, cfg):
        ctx = cfg.ctx
        pf, nf, memf, _k1 = _buf(f_vals)
        pc, nc, memc, _k2 = _buf(cp_vals)
        assert memf == memc
        h = C.c_void_p()
        ctx.check(ctx.L.smw_merkle_new_pairs(ctx.h, cfg.params.h, int(cfg.arity), int(cfg.tree_label), pf, pc, nf, nc, memf, C.byref(h)))
        return cls(ctx, h, cfg)

    @classmethod
    def new_legacy(cls, ctx, leaves, ds_tag, params=None):
        """merkle/src/lib.rs:196-223: chunks of poseidon::RATE = 16, every node hash_with_ds(chunk, ds_tag); params None = merkle::default_params"""
        p, n, mem, _k = _buf(leaves)
        tag = _fr1(ds_tag)
        h = C.c_void_p()
        ctx.check(ctx.L.smw_merkle_new_legacy(ctx.h, params.h if params else None, C.c_void_p(tag.ctypes.data), p, n, mem, C.byref(h)))
        return cls(ctx, h, None)

    @classmethod
    def new_pairs_legacy(cls, ctx, f_vals, cp_vals, ds_tag, params=None):
        """merkle/src/lib.rs:448-470: leaf i = hash_with_ds([f_i, cp_i], ds_tag), then new_legacy"""
        pf, nf, memf, _k1 = _buf(f_vals)
        pc, nc, memc, _k2 = _buf(cp_vals)
        assert memf == memc
        tag = _fr1(ds_tag)
        h = C.c_void_p()
        ctx.check(ctx.L.smw_merkle_new_pairs_legacy(ctx.h, params.h if params else None, C.c_void_p(tag.ctypes.data), pf, pc, nf, nc, memf, C.byref(h)))
        return cls(ctx, h, None)

    def root(self):
        out = _out(1)
        self.ctx.check(self.ctx.L.smw_tree_root(self.h, C.c_void_p(out.ctypes.data)))
        return out[0]

    def arity(self):
        return int(self.ctx.L.smw_tree_arity(self.h))

    def height(self):
        return int(self.ctx.L.smw_tree_num_levels(self.h)) - 1

    def level(self, l):
        n = int(self.ctx.L.smw_tree_level_len(self.h, int(l)))
        out = _out(n)
        self.ctx.check(self.ctx.L.smw_tree_level(self.h, int(l), C.c_void_p(out.ctypes.data), HOST))
        return out

    @property
    def levels(self):
        return [self.level(l) for l in range(self.height() + 1)]

    def open_many(self, indices):
        """open_union_of_paths (merkle/src/lib.rs:246-315)"""
        idx = np.ascontiguousarray(indices, dtype=np.uint64)
        h = C.c_void_p()
        self.ctx.check(self.ctx.L.smw_tree_open_many(self.h, C.c_void_p(idx.ctypes.data) if idx.size else None, idx.size, C.byref(h)))
        return MerkleProof(self.ctx, h, self.arity())

    open_many_single = open_many

    def __del__(self):
        try:
            if self.owned and self.ctx.h:
                self.ctx.L.smw_tree_free(self.h)
        except Exception:
            pass


def merkle_commit_columns(ctx, cols, arity, params, label0=0, label_stride=1):
    """One MerkleTree::new per column, batched on the device; returns the roots (n_cols, 4)."""
    if _is_torch(cols):
        n_cols, n = cols.shape[0], cols.shape[1]
        assert cols.is_contiguous() and cols.element_size() == 8 and cols.shape[-1] == 4
        p, mem, keep = C.c_void_p(cols.data_ptr()), (DEVICE if cols.is_cuda else HOST), cols     # a (pinned) CPU tensor is a host buffer: staged copies
    else:
        keep = np.ascontiguousarray(cols, dtype=np.uint64)
        n_cols, n = keep.shape[0], keep.shape[1]
        p, mem = C.c_void_p(keep.ctypes.data), HOST
    roots = _out(n_cols)
    ctx.check(ctx.L.smw_merkle_commit_columns(ctx.h, params.h, int(arity), int(label0), int(label_stride), p, n, n_cols, mem, C.c_void_p(roots.ctypes.data)))
    return roots


# ----------------------------------------------------------------------------- commitment
class MerkleConfig:
    """commitment/src/lib.rs:31-46"""

    def __init__(self, ctx, ds_tag, params):
        self.ctx, self.ds_tag, self.params = ctx, int(ds_tag), params

    @classmethod
    def with_default_params(cls, ctx, ds_tag):
        return cls(ctx, ds_tag, commitment_default_params(ctx))


class MerkleAux:
    def __init__(self, tree):
        self.tree = tree


class MerkleCommitment:
    """CommitmentScheme for MerkleCommitment (commitment/src/lib.rs:60-113): arity 16, tree_label = ds_tag."""

    def __init__(self, cfg):
        self.cfg = cfg

    def tree_cfg(self):
        return MerkleChannelCfg(self.cfg.ctx, 16, self.cfg.params, self.cfg.ds_tag)

    def commit(self, leaves):
        ctx = self.cfg.ctx
        p, n, mem, _k = _buf(leaves)
        h = C.c_void_p()
        ctx.check(ctx.L.smw_commitment_commit(ctx.h, self.cfg.ds_tag, self.cfg.params.h, p, n, mem, C.byref(h)))
        tree = MerkleTree(ctx, h, self.tree_cfg())
        return tree.root(), MerkleAux(tree)

    def open(self, indices, aux):
        return aux.tree.open_many(indices)


# ----------------------------------------------------------------------------- NTT / LDE
def fft(ctx, coeffs):
    """fft::fft (fft/src/lib.rs:22-26): natural-order coefficients -> natural-order evaluations. (n,4) or (cols,n,4)."""
    return _ntt(ctx, coeffs, 0)


def ifft(ctx, evals):
    """fft::ifft (fft/src/lib.rs:28-32)"""
    return _ntt(ctx, evals, 1)


def _ntt(ctx, a, inverse):
    if _is_torch(a):
        n = a.shape[-2]
        cols = a.numel() // (4 * n)
        ctx.check(ctx.L.smw_ntt(ctx.h, C.c_void_p(a.data_ptr()), n.bit_length() - 1, inverse, cols, DEVICE))
        return a
    v = np.ascontiguousarray(a, dtype=np.uint64).copy()
    n = v.shape[-2]
    assert n & (n - 1) == 0
    ctx.check(ctx.L.smw_ntt(ctx.h, C.c_void_p(v.ctypes.data), n.bit_length() - 1, inverse, v.size // (4 * n), HOST))
    return v


def f52_dot(ctx, a, b, k):
    """self-test entry of the FP64-pipe arithmetic: out[i] = sum_{j<k} a[i k + j] * b[i k + j]"""
    pa, n, mem, _k1 = _buf(a)
    pb, _n, _m, _k2 = _buf(b)
    out = _out(n // k)
    ctx.check(ctx.L.smw_diag_f52_dot(ctx.h, pa, pb, n // k, int(k), C.c_void_p(out.ctypes.data), HOST))
    return out


def scale_powers(ctx, buf, g):
    """buf[i] *= g^i in place (host numpy (n, 4) uint64 or a CUDA torch tensor): twiddle step of the four-step NTT, coset scaling"""
    p, n, mem, _k = _buf(buf)
    gg = _fr1(g)
    ctx.check(ctx.L.smw_scale_powers(ctx.h, p, n, C.c_void_p(gg.ctypes.data), mem))
    return _k


def dft_strided(ctx, buf, log_g, stride, count):
    """in place: for every j < count, (buf[a * stride + j])_{a < 2^log_g} -> its DFT with the order-2^log_g root of unity"""
    p, n, mem, _k = _buf(buf)
    assert n >= ((1 << log_g) - 1) * stride + count
    ctx.check(ctx.L.smw_dft_strided(ctx.h, p, int(log_g), int(stride), int(count), mem))
    return _k


def lde(ctx, cols, log_blowup, shift, out=None):
    """coset low-degree extension. cols: (n,4) or (n_cols,n,4) -> (..., n << log_blowup, 4)"""
    sh = _fr1(shift)
    if _is_torch(cols):
        n = cols.shape[-2]
        nc = cols.numel() // (4 * n)
        assert out is not None
        ctx.check(ctx.L.smw_lde(ctx.h, C.c_void_p(cols.data_ptr()), n.bit_length() - 1, int(log_blowup), C.c_void_p(sh.ctypes.data), C.c_void_p(out.data_ptr()), nc, DEVICE))
        return out
    v = np.ascontiguousarray(cols, dtype=np.uint64)
    n = v.shape[-2]
    nc = v.size // (4 * n)
    o = np.zeros(v.shape[:-2] + (n << log_blowup, 4), dtype=np.uint64)
    ctx.check(ctx.L.smw_lde(ctx.h, C.c_void_p(v.ctypes.data), n.bit_length() - 1, int(log_blowup), C.c_void_p(sh.ctypes.data), C.c_void_p(o.ctypes.data), nc, HOST))
    return o


def lde_commit(ctx, cols, log_blowup, shift, label0=0):
    """LDE of every column, then MerkleCommitment::commit of every extended column (tree_label = label0 + column). Roots only."""
    sh = _fr1(shift)
    if _is_torch(cols):
        n_cols, n = cols.shape[0], cols.shape[1]
        assert cols.is_contiguous() and cols.element_size() == 8 and cols.shape[-1] == 4
        # a (pinned) CPU tensor is a HOST buffer: the library stages it with double-buffered copies under the kernels.  (Passed as a device
        # pointer, pinned memory still worked through unified addressing, but the first kernel then read all of it over PCIe: 47 ms per cfg2 step.)
        p, mem, keep = C.c_void_p(cols.data_ptr()), (DEVICE if cols.is_cuda else HOST), cols
    else:
        keep = np.ascontiguousarray(cols, dtype=np.uint64)
        n_cols, n = keep.shape[0], keep.shape[1]
        p, mem = C.c_void_p(keep.ctypes.data), HOST
    roots = _out(n_cols)
    ctx.check(ctx.L.smw_lde_commit(ctx.h, p, n.bit_length() - 1, int(log_blowup), C.c_void_p(sh.ctypes.data), n_cols, int(label0), mem, C.c_void_p(roots.ctypes.data)))
    return roots


# ----------------------------------------------------------------------------- transcript
class Transcript:
    """transcript/src/lib.rs:48-117; the sponge state lives on the device."""

    def __init__(self, ctx, label, params=None):
        self.ctx = ctx
        buf, n = _bytes(label)
        h = C.c_void_p()
        ctx.check(ctx.L.smw_transcript_new(ctx.h, buf, n, params.h if params else None, C.byref(h)))
        self.h = h
        self._params = params

    @classmethod
    def new(cls, ctx, label, params=None):
        return cls(ctx, label, params)

    def absorb_bytes(self, b):
        buf, n = _bytes(b)
        self.ctx.check(self.ctx.L.smw_transcript_absorb_bytes(self.h, buf, n))

    def absorb_field(self, x):
        a = _fr1(x)
        self.ctx.check(self.ctx.L.smw_transcript_absorb_field(self.h, C.c_void_p(a.ctypes.data)))

    def absorb_fields(self, xs):
        p, n, mem, _k = _buf(xs)
        self.ctx.check(self.ctx.L.smw_transcript_absorb_fields(self.h, p, n, mem))

    def challenge(self, label):
        buf, n = _bytes(label)
        out = _out(1)
        self.ctx.check(self.ctx.L.smw_transcript_challenge(self.h, buf, n, C.c_void_p(out.ctypes.data)))
        return out[0]

    def challenges(self, label, n):
        return [self.challenge(bytes(label) + int(i).to_bytes(8, "little")) for i in range(n)]

    def state(self):
        st = np.zeros((17, 4), np.uint64)
        pos = C.c_int32(0)
        self.ctx.check(self.ctx.L.smw_transcript_state(self.h, C.c_void_p(st.ctypes.data), C.byref(pos)))
        return st, pos.value

    def __del__(self):
        try:
            if self.ctx.h:
                self.ctx.L.smw_transcript_free(self.h)
        except Exception:
            pass


def tr_hash_fields_tagged(ctx, tag, fields):
    """deep_ali/src/fri.rs:28-35"""
    buf, n = _bytes(tag)
    p, nf, mem, _k = _buf(np.ascontiguousarray(fields, dtype=np.uint64).reshape(-1, 4) if not _is_torch(fields) else fields)
    out = _out(1)
    ctx.check(ctx.L.smw_tr_hash_fields_tagged(ctx.h, buf, n, p if nf else None, nf, mem, C.c_void_p(out.ctypes.data)))
    return out[0]


def hash_leaf_pair(ctx, f, s):
    """deep_ali/src/fri.rs:38-44, element-wise over arrays"""
    pf, n, mem, _k1 = _buf(f)
    ps, n2, mem2, _k2 = _buf(s)
    assert n == n2 and mem == mem2 == HOST
    out = _out(n)
    ctx.check(ctx.L.smw_hash_leaf_pair(ctx.h, pf, ps, n, C.c_void_p(out.ctypes.data), HOST))
    return out


# ----------------------------------------------------------------------------- FRI
def fri_sample_z_ell(ctx, seed_z, level, domain_size):
    """deep_ali/src/fri.rs:59-82"""
    out = _out(1)
    ctx.check(ctx.L.smw_fri_sample_z_ell(ctx.h, int(seed_z), int(level), int(domain_size), C.c_void_p(out.ctypes.data)))
    return out[0]


def fri_fold_layer(ctx, f_l, z_l, m):
    """deep_ali/src/fri.rs:85-102"""
    p, n, mem, _k = _buf(f_l)
    z = _fr1(z_l)
    out = _out(n // m if m else 0)
    ctx.check(ctx.L.smw_fri_fold_layer(ctx.h, p, n, C.c_void_p(z.ctypes.data), int(m), C.c_void_p(out.ctypes.data), HOST))
    return out


def compute_s_layer(ctx, f_l, z_l, m):
    """deep_ali/src/fri.rs:123-143"""
    p, n, mem, _k = _buf(f_l)
    z = _fr1(z_l)
    out = _out(n)
    ctx.check(ctx.L.smw_compute_s_layer(ctx.h, p, n, C.c_void_p(z.ctypes.data), int(m), C.c_void_p(out.ctypes.data), HOST))
    return out


def pick_arity_for_layer(n, requested_m):
    """deep_ali/src/fri.rs:220-229"""
    return int(_lib.load().smw_pick_arity_for_layer(int(n), int(requested_m)))


class FriProverParams:
    """deep_ali/src/fri.rs:208"""

    def __init__(self, schedule, seed_z):
        self.schedule, self.seed_z = list(schedule), int(seed_z)


class FriLayerCommitment:
    """deep_ali/src/fri.rs:194-204 (f, s and tree are fetched from the device on access)"""

    def __init__(self, st, l):
        self._st, self._l = st, l
        n, m, ar, hs = C.c_size_t(), C.c_size_t(), C.c_int32(), C.c_int32()
        st.ctx.check(st.ctx.L.smw_fri_layer_info(st.h, l, C.byref(n), C.byref(m), C.byref(ar), C.byref(hs)))
        self.n, self.m, self.arity, self.hashed_leaves = n.value, m.value, ar.value, bool(hs.value)

    @property
    def f(self):
        out = _out(self.n)
        self._st.ctx.check(self._st.ctx.L.smw_fri_layer_f(self._st.h, self._l, C.c_void_p(out.ctypes.data), HOST))
        return out

    @property
    def s(self):
        out = _out(self.n)
        self._st.ctx.check(self._st.ctx.L.smw_fri_layer_s(self._st.h, self._l, C.c_void_p(out.ctypes.data), HOST))
        return out

    @property
    def tree(self):
        ctx = self._st.ctx
        h = C.c_void_p(ctx.L.smw_fri_layer_tree(self._st.h, self._l))
        return MerkleTree(ctx, h, None, owned=False)

    @property
    def root(self):
        return self.tree.root()


class FriProverState:
    """deep_ali/src/fri.rs:210-216"""

    def __init__(self, ctx, h, schedule):
        self.ctx, self.h = ctx, h
        self.schedule = list(schedule)
        self.layers = [FriLayerCommitment(self, l) for l in range(len(schedule) + 1)]

    @property
    def z_layers(self):
        out = _out(max(len(self.schedule), 1))
        self.ctx.check(self.ctx.L.smw_fri_z(self.h, C.c_void_p(out.ctypes.data)))
        return out[: len(self.schedule)]

    def roots(self):
        out = _out(len(self.schedule) + 1)
        self.ctx.check(self.ctx.L.smw_fri_roots(self.h, C.c_void_p(out.ctypes.data)))
        return out

    @property
    def f_layers(self):
        return [l.f for l in self.layers]

    @property
    def s_layers(self):
        return [l.s for l in self.layers]

    def __del__(self):
        try:
            if self.ctx.h:
                self.ctx.L.smw_fri_free(self.h)
        except Exception:
            pass


def fri_build_transcript(ctx, f0, params, level0=0):
    """deep_ali/src/fri.rs:231-312 (the FriDomain argument only carries n0 = len(f0)).  level0 > 0: f0 is layer `level0`
    of a larger instance (tree labels and z_l level indices start there) — used by the range-sharded commit phase."""
    p, n, mem, _k = _buf(f0)
    sch = np.ascontiguousarray(params.schedule, dtype=np.uint64)
    h = C.c_void_p()
    ctx.check(ctx.L.smw_fri_build_transcript_from(ctx.h, p, n, C.c_void_p(sch.ctypes.data) if sch.size else None, len(params.schedule), params.seed_z,
                                                  int(level0), mem, C.byref(h)))
    return FriProverState(ctx, h, params.schedule)


class FriShard:
    """One rank's part of the range-sharded commit phase of a single large polynomial: local folds, leaf
    hashes and subtrees of f0[first_index, first_index + len(f0_slice)); `caps` are gathered and finished by shard.py."""

    def __init__(self, ctx, f0_slice, n0, first_index, params):
        self.ctx = ctx
        p, n, mem, _k = _buf(f0_slice)
        sch = np.ascontiguousarray(params.schedule, dtype=np.uint64)
        self.h = C.c_void_p()
        ctx.check(ctx.L.smw_fri_shard_build(ctx.h, p, n, int(n0), int(first_index), C.c_void_p(sch.ctypes.data) if sch.size else None,
                                            len(params.schedule), params.seed_z, mem, C.byref(self.h)))
        self.num_local_layers = int(ctx.L.smw_fri_shard_num_local_layers(self.h))

    def cap(self, l):
        """(nodes (k, 4) uint64, DS level of the parents of these nodes, tree arity) of layer l"""
        n, lev, ar = C.c_size_t(), C.c_int32(), C.c_int32()
        self.ctx.check(self.ctx.L.smw_fri_shard_cap_info(self.h, l, C.byref(n), C.byref(lev), C.byref(ar)))
        out = _out(n.value)
        self.ctx.check(self.ctx.L.smw_fri_shard_cap(self.h, l, C.c_void_p(out.ctypes.data), HOST))
        return out, lev.value, ar.value

    def tail(self):
        """this rank's slice of the first layer that is too small to split (empty when every layer was local)"""
        n = int(self.ctx.L.smw_fri_shard_tail_len(self.h))
        out = _out(n)
        if n:
            self.ctx.check(self.ctx.L.smw_fri_shard_tail(self.h, C.c_void_p(out.ctypes.data), HOST))
        return out

    def __del__(self):
        try:
            if self.ctx.h and self.h:
                self.ctx.L.smw_fri_shard_free(self.h)
        except Exception:
            pass


def merkle_root_from_level(ctx, nodes, level, arity, params, tree_label):
    """root of the tree whose level-`level` nodes are `nodes` (level 0 = leaves / leaf digests), merkle/src/lib.rs:164-181"""
    p, n, mem, _k = _buf(nodes)
    out = _out(1)
    ctx.check(ctx.L.smw_merkle_root_from_level(ctx.h, params.h, int(arity), int(tree_label), int(level), p, n, mem, C.c_void_p(out.ctypes.data)))
    return out[0]


# ----------------------------------------------------------------------------- Merkle-folded sum-check
def sumcheck_round_coeffs(ctx, layer):
    """channel/src/lib.rs:406-416"""
    p, n, mem, _k = _buf(layer)
    c0, c1 = _out(1), _out(1)
    ctx.check(ctx.L.smw_sumcheck_round_coeffs(ctx.h, p, n, C.c_void_p(c0.ctypes.data), C.c_void_p(c1.ctypes.data), mem))
    return c0[0], c1[0]


def sumcheck_round_coeffs_noalloc(ctx, layer_dev, out_dev):
    """smw_sumcheck_round_coeffs_noalloc: no allocation anywhere (fixed 64 KiB scratch inside the library, n <= 2^26).  `layer_dev` is a
    CUDA tensor of n elements (n, 4) int64/uint64, `out_dev` a CUDA tensor of 8 x 8 bytes that receives c0 then c1; asynchronous on the
    context's stream.  No counterpart in the reference."""
    assert layer_dev.is_cuda and out_dev.is_cuda and layer_dev.is_contiguous() and out_dev.is_contiguous()
    assert out_dev.numel() * out_dev.element_size() >= 64
    n = layer_dev.numel() * layer_dev.element_size() // 32
    ctx.check(ctx.L.smw_sumcheck_round_coeffs_noalloc(ctx.h, C.c_void_p(layer_dev.data_ptr()), n, C.c_void_p(out_dev.data_ptr())))


def mf_round_challenge_from_root(ctx, round_idx, prev_root, tr_params=None):
    """channel/src/lib.rs:592-598"""
    r = _fr1(prev_root)
    out = _out(1)
    ctx.check(ctx.L.smw_mf_round_challenge_from_root(ctx.h, int(round_idx), C.c_void_p(r.ctypes.data), tr_params.h if tr_params else None, C.c_void_p(out.ctypes.data)))
    return out[0]


def mf_fold(ctx, cur, r):
    """channel/src/lib.rs:648-655"""
    p, n, mem, _k = _buf(cur)
    rr = _fr1(r)
    out = _out(n // 2)
    ctx.check(ctx.L.smw_mf_fold(ctx.h, p, n, C.c_void_p(rr.ctypes.data), C.c_void_p(out.ctypes.data), HOST))
    return out


def mf_commit_phase(ctx, table, ds_tag):
    """SumCheckMFProver::new + the commit half of every round (channel/src/lib.rs:600-657): roots, r_i, final value"""
    p, n, mem, _k = _buf(table)
    k = n.bit_length() - 1
    roots, rs, fin = _out(k + 1), _out(max(k, 1)), _out(1)
    ctx.check(ctx.L.smw_mf_commit_phase(ctx.h, int(ds_tag), p, n, mem, C.c_void_p(roots.ctypes.data), C.c_void_p(rs.ctypes.data), C.c_void_p(fin.ctypes.data)))
    return roots, rs[:k], fin[0]


def deep_ali_merge_evals(ctx, a, s, e, t, omega, z):
    """deep_ali/src/lib.rs:47-105, f0 only"""
    pa, n, mem, _k1 = _buf(a)
    ps, _, _, _k2 = _buf(s)
    pe, _, _, _k3 = _buf(e)
    pt, _, _, _k4 = _buf(t)
    w, zz = _fr1(omega), _fr1(z)
    out = _out(n)
    ctx.check(ctx.L.smw_deep_ali_merge_evals(ctx.h, pa, ps, pe, pt, n, C.c_void_p(w.ctypes.data), C.c_void_p(zz.ctypes.data), C.c_void_p(out.ctypes.data), HOST))
    return out


# ----------------------------------------------------------------------------- query phase + DEEP-FRI prover ("next" tier)
def _fr_from_u64(ctx, v):
    c = np.zeros((1, 4), dtype=np.uint64)
    c[0, 0] = int(v)
    return ctx.from_canonical(c)[0]


def index_from_seed(ctx, seeds, n_pow2):
    """deep_ali/src/fri.rs:178-183, element-wise: StdRng::from_seed(canonical bytes).next_u64() & (n_pow2 - 1)"""
    p, n, mem, _k = _buf(seeds)
    out = np.zeros(max(n, 1), dtype=np.uint64)
    ctx.check(ctx.L.smw_index_from_seed(ctx.h, p, n, int(n_pow2), C.c_void_p(out.ctypes.data), mem))
    return [int(v) for v in out[:n]]


def ali_sample_z_beta_fs(ctx, tag, n0, roots_seed):
    """deep_ali/src/fri.rs:508-533 -> (z, beta)"""
    fused = tr_hash_fields_tagged(ctx, tag, np.stack([np.asarray(roots_seed, dtype=np.uint64).reshape(4), _fr_from_u64(ctx, n0)]))
    z, beta, rs = _out(1), _out(1), _fr1(roots_seed)
    ctx.check(ctx.L.smw_ali_sample_z_beta(ctx.h, C.c_void_p(fused.ctypes.data), C.c_void_p(rs.ctypes.data), int(n0),
                                          C.c_void_p(z.ctypes.data), C.c_void_p(beta.ctypes.data)))
    return z[0], beta[0]


def tr_hash_fields_tagged_many(ctx, tags, xss):
    """[tr_hash_fields_tagged(tag_i, xs_i)] (deep_ali/src/fri.rs:28-36) with the serial sponges overlapped on the device"""
    k = len(tags)
    bufs = [_buf(x) for x in xss]
    assert all(b[2] == bufs[0][2] for b in bufs)
    tag_bufs = [(C.c_uint8 * max(len(t), 1)).from_buffer_copy(bytes(t) if len(t) else b"\0") for t in tags]
    tag_ptrs = (C.c_void_p * k)(*[C.cast(tb, C.c_void_p) for tb in tag_bufs])
    tag_lens = (C.c_size_t * k)(*[len(t) for t in tags])
    x_ptrs = (C.c_void_p * k)(*[b[0] for b in bufs])
    ns = (C.c_size_t * k)(*[b[1] for b in bufs])
    out = _out(k)
    ctx.check(ctx.L.smw_tr_hash_fields_tagged_many(ctx.h, k, tag_ptrs, tag_lens, x_ptrs, ns, bufs[0][2], C.c_void_p(out.ctypes.data)))
    return out


def deep_ali_build_f0(ctx, a, s, e, t):
    """DeepAliRealBuilder::default().build_f0 (deep_ali/src/fri.rs:535-570): no blinding, tag "ALI/DEEP" -> (f0, z)"""
    n0 = len(a)
    h = list(tr_hash_fields_tagged_many(ctx, [b"ALI/A", b"ALI/S", b"ALI/E", b"ALI/T"], [a, s, e, t]))      # four serial sponges, side by side
    seed_f = tr_hash_fields_tagged(ctx, b"ALI/seed", np.stack(h + [_fr_from_u64(ctx, n0)]))
    z, _beta = ali_sample_z_beta_fs(ctx, b"ALI/DEEP", n0, seed_f)
    omega = ctx.root_of_unity(n0.bit_length() - 1)                      # FriDomain::new_radix2 (fri.rs:53-56)
    return deep_ali_merge_evals(ctx, a, s, e, t, omega, z), z


class FriQueryProofs:
    """what fri_prove_queries (deep_ali/src/fri.rs:355-466) hands to the verifier, as far as proof size goes: the roots seed,
    per layer the sampled child / parent indices and the two batched multiproofs, and the opening of the final layer"""

    def __init__(self, roots_seed, child_idx, parent_idx, child_proofs, parent_proofs, final_proof, r, L):
        self.roots_seed, self.child_idx, self.parent_idx = roots_seed, child_idx, parent_idx
        self.child_proofs, self.parent_proofs, self.final_proof, self.r, self.L = child_proofs, parent_proofs, final_proof, r, L

    def size_bytes(self):
        """deep_fri_proof_size_bytes (deep_ali/src/fri.rs:775-805)"""
        nsib = lambda pr: sum(len(x) for x in pr.siblings)       # noqa: E731
        total = (self.L + 1) * 32 + 32 + 8
        for c, p in zip(self.child_proofs, self.parent_proofs):
            total += (nsib(c) + nsib(p)) * 32 + (len(c.indices) + len(p.indices)) * 8
        total += nsib(self.final_proof) * 32
        total += self.r * (8 + 64 + self.L * 16 + self.L * 128)
        return total


def tr_hash_fields_tagged_batch(ctx, tag, xs):
    """xs: (count, k, 4) -> (count, 4): count one-shot hashes tr_hash_fields_tagged(tag, xs[i]) in one launch (k small)"""
    a = np.ascontiguousarray(xs, dtype=np.uint64)
    count, k = a.shape[0], a.shape[1]
    buf, n = _bytes(tag)
    out = _out(count)
    ctx.check(ctx.L.smw_tr_hash_fields_tagged_batch(ctx.h, buf, n, C.c_void_p(a.ctypes.data), k, count, HOST, C.c_void_p(out.ctypes.data)))
    return out


def fri_prove_queries(ctx, st, r):
    """deep_ali/src/fri.rs:355-466: query indices from the roots (fs_seed_from_roots :176, index_seed :184-191 with its retry
    rule) and the per-layer batched openings; every hash, RNG draw and opening runs on the device."""
    L = len(st.schedule)
    roots = st.roots()
    roots_seed = tr_hash_fields_tagged(ctx, b"FRI/seed", roots)
    one = _fr_from_u64(ctx, 1)
    child_idx, parent_idx, cps, pps = [], [], [], []
    # every index seed H("FRI/index", [roots_seed, l, q]) of the proof in ONE launch (r * L hashes, one warp each)
    small = ctx.from_canonical(np.array([[v, 0, 0, 0] for v in range(max(L, r, 1))], dtype=np.uint64))      # the integers 0 .. as field elements
    pre = np.empty((L, r, 3, 4), dtype=np.uint64)
    pre[:, :, 0] = roots_seed
    pre[:, :, 1] = small[:L, None, :] if L else 0
    pre[:, :, 2] = small[None, :r, :]
    all_seeds = tr_hash_fields_tagged_batch(ctx, b"FRI/index", pre.reshape(L * r, 3, 4)).reshape(L, r, 4) if L * r else np.zeros((L, r, 4), dtype=np.uint64)
    for l in range(L):
        n, m = st.layers[l].n, st.layers[l].m
        np2 = 1 << max(n - 1, 0).bit_length()
        seeds = all_seeds[l]
        idx = index_from_seed(ctx, seeds, np2)
        for q in range(r):
            if idx[q] >= n:                                             # fri.rs:186-190: one re-hash, then wrap
                rs = tr_hash_fields_tagged(ctx, b"FRI/index", np.stack([seeds[q], one]))
                i2 = index_from_seed(ctx, rs.reshape(1, 4), np2)[0]
                idx[q] = i2 if i2 < n else (i2 & (n - 1))
        ci, pi = idx, [i // m for i in idx]
        cps.append(st.layers[l].tree.open_many(ci))
        pps.append(st.layers[l + 1].tree.open_many(pi))
        child_idx.append(ci)
        parent_idx.append(pi)
    final = st.layers[L].tree.open_many([0])
    return FriQueryProofs(roots_seed, child_idx, parent_idx, cps, pps, final, r, L)


def deep_fri_prove(ctx, a, s, e, t, schedule, r, seed_z):
    """deep_ali/src/fri.rs:601-641 -> (FriProverState, FriQueryProofs); .size_bytes() is the reference's published proof_bytes"""
    f0, _z = deep_ali_build_f0(ctx, a, s, e, t)
    st = fri_build_transcript(ctx, f0, FriProverParams(schedule, seed_z))
    return st, fri_prove_queries(ctx, st, r)


# ------------------------------------------------------------------------------------------------ DeepFriProof + verifier
class DeepFriParams:
    """deep_ali/src/fri.rs:572-577"""

    def __init__(self, schedule, r, seed_z):
        self.schedule, self.r, self.seed_z = [int(m) for m in schedule], int(r), int(seed_z)


class LayerBatchProof:
    """deep_ali/src/fri.rs:316-326"""

    def __init__(self, hashed_leaves, child_indices, child_proof, parent_indices, parent_proof):
        self.hashed_leaves = bool(hashed_leaves)
        self.child_indices, self.child_proof = [int(i) for i in child_indices], child_proof
        self.parent_indices, self.parent_proof = [int(i) for i in parent_indices], parent_proof


class FriQueryPayload:
    """deep_ali/src/fri.rs:579-594: per layer (i, parent_index) and (f_i, s_i, f_parent_b, s_parent_b); the final pair"""

    def __init__(self, refs, payloads, final_index, final_pair):
        self.per_layer_refs, self.per_layer_payloads = refs, payloads       # [(i, b)], [(4, 4) uint64: f_i, s_i, f_parent_b, s_parent_b]
        self.final_index, self.final_pair = int(final_index), final_pair


class DeepFriProof:
    """deep_ali/src/fri.rs:596-602"""

    def __init__(self, roots, layers, final_proof, queries, n0, omega0):
        self.roots, self.layers, self.final_proof, self.queries, self.n0, self.omega0 = roots, layers, final_proof, queries, int(n0), omega0


def deep_fri_proof_from_state(ctx, st, qp, params):
    """the assembly half of deep_fri_prove (deep_ali/src/fri.rs:618-640): batches + the per-query field payloads.  The opened
    values are gathered from the layers the prover state keeps on the device."""
    L = len(params.schedule)
    fs = [(st.layers[l].f, st.layers[l].s) for l in range(L + 1)]
    layers = []
    for l in range(L):
        layers.append(LayerBatchProof(st.layers[l].hashed_leaves, qp.child_proofs[l].indices, qp.child_proofs[l], qp.parent_proofs[l].indices, qp.parent_proofs[l]))
    queries = []
    for q in range(params.r):
        refs, pays = [], []
        for l in range(L):
            i, b = int(qp.child_idx[l][q]), int(qp.parent_idx[l][q])
            refs.append((i, b))
            pays.append(np.stack([fs[l][0][i], fs[l][1][i], fs[l + 1][0][b], fs[l + 1][1][b]]))
        queries.append(FriQueryPayload(refs, pays, 0, (fs[L][0][0].copy(), fs[L][1][0].copy())))
    return DeepFriProof(st.roots(), layers, qp.final_proof, queries, st.layers[0].n, ctx.root_of_unity(st.layers[0].n.bit_length() - 1))


def deep_fri_prove_proof(ctx, a, s, e, t, params):
    """deep_fri_prove (deep_ali/src/fri.rs:604-641) with DeepAliRealBuilder::default() -> DeepFriProof"""
    st, qp = deep_fri_prove(ctx, a, s, e, t, params.schedule, params.r, params.seed_z)
    return deep_fri_proof_from_state(ctx, st, qp, params)


def _layer_is_hashed(arity):
    return arity in (128, 64, 32, 16, 8)


def _verify_layer_opening(ctx, root, indices, pair_of, proof, n_layer, m, label):
    """one batched opening of deep_fri_verify (fri.rs:672-727): hashed layers go through hash_leaf_pair + verify_single,
    the others through verify_pairs; `pair_of` maps index -> (f, s) or is missing it (-> False)"""
    arity = pick_arity_for_layer(n_layer, m)
    gp = poseidon_params_for_arity(ctx, arity)             # MerkleChannelCfg::new(arity)
    vals = []
    for i in indices:
        if i not in pair_of:
            return False
        vals.append(pair_of[i])
    if not vals:
        return verify_many_ds(root, [], np.zeros((0, 4), np.uint64), proof, label, gp)
    pairs = np.stack([np.stack(v) for v in vals])          # (k, 2, 4)
    if _layer_is_hashed(arity):
        leaves = hash_leaf_pair(ctx, np.ascontiguousarray(pairs[:, 0]), np.ascontiguousarray(pairs[:, 1]))
        return verify_many_ds(root, indices, leaves, proof, label, gp)
    return verify_pairs_ds(root, indices, pairs, proof, label, gp)


def deep_fri_verify(ctx, params, proof):
    """deep_ali/src/fri.rs:643-762, check for check: structure, the child / parent multiproofs of every layer against the
    roots, the local fold relation s_i == f_parent[b], the opening of the final layer at index 0.  Hashes run on the device."""
    L = len(params.schedule)
    if len(proof.roots) != L + 1 or len(proof.layers) != L or len(proof.queries) != params.r:
        return False
    sizes = [proof.n0]
    for m in params.schedule:
        if sizes[-1] % m:
            raise SmwError(5, "schedule not dividing domain size")          # layer_sizes_from_schedule asserts (fri.rs:150)
        sizes.append(sizes[-1] // m)
    child_maps, parent_maps = [dict() for _ in range(L)], [dict() for _ in range(L)]
    for qp in proof.queries:
        if len(qp.per_layer_refs) != L or len(qp.per_layer_payloads) != L:
            return False
        for l in range(L):
            (i, b), pay = qp.per_layer_refs[l], qp.per_layer_payloads[l]
            child_maps[l].setdefault(int(i), (pay[0], pay[1]))               # entry().or_insert: the first payload wins
            parent_maps[l].setdefault(int(b), (pay[2], pay[3]))
    for l in range(L):
        lb = proof.layers[l]
        if not _verify_layer_opening(ctx, proof.roots[l], lb.child_indices, child_maps[l], lb.child_proof, sizes[l], params.schedule[l], l):
            return False
        m_next = params.schedule[l + 1] if l + 1 < L else 1
        if not _verify_layer_opening(ctx, proof.roots[l + 1], lb.parent_indices, parent_maps[l], lb.parent_proof, sizes[l + 1], m_next, l + 1):
            return False
    for qp in proof.queries:                                                   # verify_local_check_fold (fri.rs:169-176)
        for l in range(L):
            (i, _b), pay = qp.per_layer_refs[l], qp.per_layer_payloads[l]
            m = params.schedule[l]
            if i // m >= sizes[l] // m or not np.array_equal(pay[1], pay[2]):
                return False
    if proof.queries[0].final_index != 0:
        return False
    fin = proof.queries[0].final_pair
    return _verify_layer_opening(ctx, proof.roots[L], [0], {0: (fin[0], fin[1])}, proof.final_proof, sizes[L], 1, L)


# ---------------------------------------------------------------------------------------------- multi-GPU data plane
class Comm:
    """smw_comm: one rank of an NCCL communicator owned by the library (csrc/comm.cu).  Not in the reference (it has no distributed
    path).  Bootstrap: `Comm.unique_id()` on rank 0, carried to every rank by the host (see shard.make_comm), `Comm(ctx, id, world, rank)`."""

    def __init__(self, ctx, uid, world, rank):
        self.ctx = ctx
        buf = (C.c_uint8 * 128).from_buffer_copy(bytes(uid))
        h = C.c_void_p()
        ctx.check(ctx.L.smw_comm_init(ctx.h, buf, int(world), int(rank), C.byref(h)))
        self.h, self.world, self.rank = h, int(world), int(rank)

    @staticmethod
    def unique_id():
        L = _lib.load()
        buf = (C.c_uint8 * 128)()
        rc = L.smw_comm_unique_id(buf)
        if rc:
            raise SmwError(rc, "ncclGetUniqueId (is libnccl.so.2 loadable?)")
        return bytes(buf)

    def all_gather(self, send):
        """send: CUDA tensor (k, 4) int64 -> (world * k, 4) on the device, rank order; on the context's stream"""
        import torch
        p, n, mem, _k = _buf(send)
        assert mem == DEVICE
        out = torch.empty((self.world * send.shape[0], 4), dtype=send.dtype, device=send.device)
        self.ctx.check(self.ctx.L.smw_comm_all_gather(self.h, p, C.c_void_p(out.data_ptr()), n))
        return out

    def close(self):
        if self.h and self.ctx.h:
            self.ctx.L.smw_comm_free(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def fri_roots_sharded(ctx, comm, chunks, first_indices, n0, params):
    """Layer roots of fri_build_transcript (deep_ali/src/fri.rs:231-312) for one polynomial held as world * len(first_indices) equal
    contiguous slices; chunks: (C, n_chunk, 4) host array or CUDA tensor; comm None = one rank.  One all-gather per proof."""
    p, n, mem, _k = _buf(chunks)
    fi = np.ascontiguousarray(first_indices, dtype=np.uint64)
    nc = int(fi.size)
    sch = np.ascontiguousarray(params.schedule, dtype=np.uint64)
    out = _out(len(params.schedule) + 1)
    ctx.check(ctx.L.smw_fri_roots_sharded(ctx.h, comm.h if comm is not None else None, p, n // nc, nc, C.c_void_p(fi.ctypes.data), int(n0),
                                          C.c_void_p(sch.ctypes.data) if sch.size else None, len(params.schedule), params.seed_z, mem, C.c_void_p(out.ctypes.data)))
    return out


def ntt_cyclic_to_chunks(ctx, comm, x_local, log_n):
    """One NTT of 2^log_n elements spread cyclically over the ranks (x_local: CUDA tensor (M, 4), overwritten) -> CUDA tensor
    (world, M / world, 4): chunk k1 holds X[M k1 + rank M / world ..]; one all-to-all (smw_ntt_cyclic_to_chunks)"""
    import torch
    world = comm.world if comm is not None else 1
    M = x_local.shape[0]
    out = torch.empty_like(x_local)
    ctx.check(ctx.L.smw_ntt_cyclic_to_chunks(ctx.h, comm.h if comm is not None else None, C.c_void_p(x_local.data_ptr()), C.c_void_p(out.data_ptr()), int(log_n)))
    return out.view(world, M // world, 4)


def mf_commit_phase_sharded(ctx, comm, table_slice, ds_tag):
    """smw_mf_commit_phase for a table held as `world` contiguous slices: (roots, r, final), identical on every rank"""
    p, n, mem, _k = _buf(table_slice)
    world = comm.world if comm is not None else 1
    k = (n * world).bit_length() - 1
    roots, rs, fin = _out(k + 1), _out(max(k, 1)), _out(1)
    ctx.check(ctx.L.smw_mf_commit_phase_sharded(ctx.h, comm.h if comm is not None else None, int(ds_tag), p, n, mem, C.c_void_p(roots.ctypes.data),
                                                C.c_void_p(rs.ctypes.data), C.c_void_p(fin.ctypes.data)))
    return roots, rs[:k], fin[0]
