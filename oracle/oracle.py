"""ctypes wrapper over oracle/liboracle.so — the CPU restatement of the reference.

TEST INFRASTRUCTURE ONLY (see the header of oracle.c): imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
Nothing under stark_mlwe_b200/ may import this module.

Field elements travel as numpy uint64 arrays of shape (..., 4): little-endian
64-bit limbs in Montgomery form (R = 2^256), byte-identical to ark-ff's Fp256.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")

P = 0x40000000000000000000000000000000224698FC0994A8DD8C46EB2100000001
R = (1 << 256) % P
R_INV = pow(R, -1, P)


def build(force=False):
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle.so"], stderr=subprocess.DEVNULL)
    return _SO


_lib = None
u64p = C.POINTER(C.c_uint64)
u8p = C.POINTER(C.c_uint8)
szp = C.POINTER(C.c_size_t)


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        L = _lib
        for name in ("orc_rng_seed_from_u64", "orc_rng_from_seed", "orc_params_new", "orc_params_for_width",
                     "orc_params_for_arity", "orc_transcript_default_params", "orc_commitment_default_params",
                     "orc_tr_new", "orc_merkle_new", "orc_merkle_new_pairs", "orc_tree_open", "orc_fri_build",
                     "orc_fri_layer_tree", "orc_commitment_commit", "orc_merkle_new_legacy", "orc_merkle_new_pairs_legacy"):
            getattr(L, name).restype = C.c_void_p
        for name in ("orc_rng_next_u64",):
            getattr(L, name).restype = C.c_uint64
        for name in ("orc_tree_level_len", "orc_proof_total_siblings", "orc_proof_nsib", "orc_proof_ngrp",
                     "orc_proof_nidx", "orc_fri_layer_n", "orc_fri_prove_size", "orc_deep_fri_prove_size"):
            getattr(L, name).restype = C.c_size_t
        L.orc_init()
    return _lib


# ----------------------------------------------------------------- helpers
def _p(a):
    return a.ctypes.data_as(u64p)


def _vp(h):
    return C.c_void_p(h)


def _bytes(b):
    b = bytes(b)
    return (C.c_uint8 * max(len(b), 1)).from_buffer_copy(b if b else b"\0"), len(b)


def fr_array(x):
    a = np.ascontiguousarray(x, dtype=np.uint64)
    assert a.shape[-1] == 4
    return a


def ints_to_limbs(vals):
    """canonical python ints -> (n,4) uint64 (no Montgomery conversion)"""
    out = np.zeros((len(vals), 4), dtype=np.uint64)
    for i, v in enumerate(vals):
        v = int(v)
        for k in range(4):
            out[i, k] = (v >> (64 * k)) & 0xFFFFFFFFFFFFFFFF
    return out


def limbs_to_ints(a):
    a = np.asarray(a, dtype=np.uint64).reshape(-1, 4)
    return [int(r[0]) | (int(r[1]) << 64) | (int(r[2]) << 128) | (int(r[3]) << 192) for r in a]


def to_mont(vals):
    """canonical ints -> Montgomery-form (n,4) array"""
    return ints_to_limbs([(int(v) % P) * R % P for v in vals])


def from_mont(a):
    """Montgomery-form array -> canonical ints"""
    return [v * R_INV % P for v in limbs_to_ints(a)]


def fr_hex(a):
    """Montgomery-form single element -> canonical hex (as SURVEY Appendix B prints)"""
    return "0x%064x" % from_mont(np.asarray(a).reshape(1, 4))[0]


def set_threads(n):
    lib().orc_set_threads(int(n))


def max_threads():
    return int(lib().orc_max_threads())


# ------------------------------------------------------------------- field
def fr_mul(a, b):
    a, b = fr_array(a), fr_array(b)
    out = np.empty_like(a)
    lib().orc_fr_mul(_p(a), _p(b), _p(out), C.c_size_t(a.size // 4))
    return out


def fr_add(a, b):
    a, b = fr_array(a), fr_array(b)
    out = np.empty_like(a)
    lib().orc_fr_add(_p(a), _p(b), _p(out), C.c_size_t(a.size // 4))
    return out


def fr_sub(a, b):
    a, b = fr_array(a), fr_array(b)
    out = np.empty_like(a)
    lib().orc_fr_sub(_p(a), _p(b), _p(out), C.c_size_t(a.size // 4))
    return out


def fr_inv(a):
    a = fr_array(a)
    out = np.empty_like(a)
    lib().orc_fr_inv(_p(a), _p(out), C.c_size_t(a.size // 4))
    return out


def to_mont_c(canon):
    canon = fr_array(canon)
    out = np.empty_like(canon)
    lib().orc_to_mont(_p(canon), _p(out), C.c_size_t(canon.size // 4))
    return out


def from_mont_c(m):
    m = fr_array(m)
    out = np.empty_like(m)
    lib().orc_from_mont(_p(m), _p(out), C.c_size_t(m.size // 4))
    return out


def root_of_unity(log_n):
    out = np.zeros(4, dtype=np.uint64)
    lib().orc_root_of_unity(int(log_n), _p(out))
    return out


def blake3(data):
    buf, n = _bytes(data)
    out = (C.c_uint8 * 32)()
    lib().orc_blake3(buf, C.c_size_t(n), out)
    return bytes(out)


def fr_from_hash(tag, data):
    buf, n = _bytes(data)
    out = np.zeros(4, dtype=np.uint64)
    lib().orc_fr_from_hash(tag.encode(), buf, C.c_size_t(n), _p(out))
    return out


# --------------------------------------------------------------------- rng
class StdRng:
    """rand 0.8.5 StdRng (ChaCha12)."""

    def __init__(self, h):
        self.h = h

    @classmethod
    def seed_from_u64(cls, s):
        return cls(lib().orc_rng_seed_from_u64(C.c_uint64(s & 0xFFFFFFFFFFFFFFFF)))

    @classmethod
    def from_seed(cls, seed32):
        buf, n = _bytes(seed32)
        assert n == 32
        return cls(lib().orc_rng_from_seed(buf))

    def next_u64(self):
        return int(lib().orc_rng_next_u64(_vp(self.h)))

    def fr_rand(self, n):
        """n draws of ark-ff `F::rand` -> Montgomery-form (n,4)"""
        out = np.empty((n, 4), dtype=np.uint64)
        lib().orc_rng_fr_rand(_vp(self.h), _p(out), C.c_size_t(n))
        return out

    def __del__(self):
        try:
            lib().orc_rng_free(_vp(self.h))
        except Exception:
            pass


# ---------------------------------------------------------------- poseidon
class Params:
    """PoseidonParamsDynamic (poseidon/src/lib.rs:104-114)."""

    def __init__(self, h, owned=True):
        self.h = h
        self.owned = owned
        self.t = int(lib().orc_params_t(_vp(h)))
        self.rp = int(lib().orc_params_rp(_vp(h)))

    @classmethod
    def from_seed(cls, seed, t=17, rp=64):
        buf, n = _bytes(seed)
        return cls(lib().orc_params_new(buf, C.c_size_t(n), t, rp))

    @classmethod
    def for_width(cls, t):
        h = lib().orc_params_for_width(int(t))
        if not h:
            raise ValueError("unsupported Poseidon width t=%d" % t)
        return cls(h)

    @classmethod
    def for_arity(cls, arity):
        h = lib().orc_params_for_arity(int(arity))
        if not h:
            raise ValueError("unsupported Merkle arity %d" % arity)
        return cls(h)

    @classmethod
    def transcript_default(cls):
        return cls(lib().orc_transcript_default_params(), owned=False)

    @classmethod
    def commitment_default(cls):
        return cls(lib().orc_commitment_default_params(), owned=False)

    def export(self):
        t = self.t
        mds = np.empty((t, t, 4), dtype=np.uint64)
        rcf = np.empty((8, t, 4), dtype=np.uint64)
        rcp = np.empty((self.rp, 4), dtype=np.uint64)
        lib().orc_params_export(_vp(self.h), _p(mds), _p(rcf), _p(rcp))
        return mds, rcf, rcp

    def __del__(self):
        if self.owned:
            try:
                lib().orc_params_free(_vp(self.h))
            except Exception:
                pass


def permute(params, states):
    """states: (batch, t, 4) Montgomery; returns permuted copy"""
    s = fr_array(states).copy().reshape(-1, params.t, 4)
    lib().orc_permute(_vp(params.h), _p(s), C.c_size_t(s.shape[0]))
    return s


def hash_with_ds_dynamic(params, ds, inputs):
    ds = fr_array(ds).reshape(-1, 4)
    inp = fr_array(inputs).reshape(-1, 4)
    out = np.zeros(4, dtype=np.uint64)
    lib().orc_hash_with_ds_dynamic(_vp(params.h), _p(ds), int(ds.shape[0]), _p(inp), C.c_size_t(inp.shape[0]), _p(out))
    return out


def hash_with_ds_legacy(params, inputs, ds_tag):
    inp = fr_array(inputs).reshape(-1, 4)
    tag = fr_array(ds_tag).reshape(4)
    out = np.zeros(4, dtype=np.uint64)
    lib().orc_hash_with_ds_legacy(_vp(params.h), _p(inp), C.c_size_t(inp.shape[0]), _p(tag), _p(out))
    return out


# -------------------------------------------------------------- transcript
class Transcript:
    """transcript/src/lib.rs:48-117"""

    def __init__(self, label, params=None):
        buf, n = _bytes(label)
        self._params = params
        self.h = lib().orc_tr_new(buf, C.c_size_t(n), _vp(params.h) if params else None)

    def absorb_bytes(self, b):
        buf, n = _bytes(b)
        lib().orc_tr_absorb_bytes(_vp(self.h), buf, C.c_size_t(n))

    def absorb_field(self, x):
        x = fr_array(x).reshape(4)
        lib().orc_tr_absorb_field(_vp(self.h), _p(x))

    def absorb_fields(self, xs):
        for x in fr_array(xs).reshape(-1, 4):
            self.absorb_field(x)

    def challenge(self, label):
        buf, n = _bytes(label)
        out = np.zeros(4, dtype=np.uint64)
        lib().orc_tr_challenge(_vp(self.h), buf, C.c_size_t(n), _p(out))
        return out

    def state(self):
        st = np.zeros((17, 4), dtype=np.uint64)
        pos = C.c_int(0)
        lib().orc_tr_state(_vp(self.h), _p(st), C.byref(pos))
        return st, pos.value

    def __del__(self):
        try:
            lib().orc_tr_free(_vp(self.h))
        except Exception:
            pass


def tr_hash_fields_tagged(tag, xs):
    xs = fr_array(xs).reshape(-1, 4)
    out = np.zeros(4, dtype=np.uint64)
    lib().orc_tr_hash_fields_tagged(bytes(tag), _p(xs), C.c_size_t(xs.shape[0]), _p(out))
    return out


def hash_leaf_pair(f, s):
    f, s = fr_array(f).reshape(-1, 4), fr_array(s).reshape(-1, 4)
    out = np.empty_like(f)
    lib().orc_hash_leaf_pair(_p(f), _p(s), _p(out), C.c_size_t(f.shape[0]))
    return out


# ------------------------------------------------------------------ merkle
class MerkleProof:
    def __init__(self, h):
        L = lib()
        self.nlev = int(L.orc_proof_nlevels(_vp(h)))
        n = int(L.orc_proof_nidx(_vp(h)))
        idx = np.zeros(max(n, 1), dtype=np.uint64)
        L.orc_proof_indices(_vp(h), _p(idx))
        self.indices = [int(v) for v in idx[:n]]
        self.siblings, self.group_sizes = [], []
        for l in range(self.nlev):
            ns, ng = int(L.orc_proof_nsib(_vp(h), l)), int(L.orc_proof_ngrp(_vp(h), l))
            sib = np.zeros((max(ns, 1), 4), dtype=np.uint64)
            grp = np.zeros(max(ng, 1), dtype=np.uint8)
            L.orc_proof_level(_vp(h), l, _p(sib), grp.ctypes.data_as(u8p))
            self.siblings.append(sib[:ns].copy())
            self.group_sizes.append(grp[:ng].copy())
        L.orc_proof_free(_vp(h))


class MerkleTree:
    """merkle/src/lib.rs:114-128; level 0 = leaves"""

    def __init__(self, h, params, arity, tree_label, owned=True):
        if not h:
            raise ValueError("merkle: empty leaves or arity/width mismatch")
        self.h, self.params, self.arity, self.tree_label, self.owned = h, params, arity, tree_label, owned

    @classmethod
    def new(cls, leaves, arity, tree_label=0, params=None):
        params = params or Params.for_arity(arity)
        lv = fr_array(leaves).reshape(-1, 4)
        return cls(lib().orc_merkle_new(_vp(params.h), int(arity), C.c_uint64(tree_label), _p(lv), C.c_size_t(lv.shape[0])), params, arity, tree_label)

    @classmethod
    def new_pairs(cls, f, cp, arity, tree_label=0, params=None):
        params = params or Params.for_arity(arity)
        f, cp = fr_array(f).reshape(-1, 4), fr_array(cp).reshape(-1, 4)
        assert f.shape == cp.shape, "f and cp length mismatch"
        return cls(lib().orc_merkle_new_pairs(_vp(params.h), int(arity), C.c_uint64(tree_label), _p(f), _p(cp), C.c_size_t(f.shape[0])), params, arity, tree_label)

    @classmethod
    def new_legacy(cls, leaves, ds_tag, params=None):
        """merkle/src/lib.rs:196-223 (fixed t = 17, chunks of 16, one ds_tag)"""
        params = params or Params.commitment_default()
        lv, tag = fr_array(leaves).reshape(-1, 4), fr_array(ds_tag).reshape(4)
        return cls(lib().orc_merkle_new_legacy(_vp(params.h), _p(tag), _p(lv), C.c_size_t(lv.shape[0])), params, 16, 0)

    @classmethod
    def new_pairs_legacy(cls, f, cp, ds_tag, params=None):
        """merkle/src/lib.rs:448-470"""
        params = params or Params.commitment_default()
        f, cp, tag = fr_array(f).reshape(-1, 4), fr_array(cp).reshape(-1, 4), fr_array(ds_tag).reshape(4)
        assert f.shape == cp.shape, "f and cp length mismatch"
        return cls(lib().orc_merkle_new_pairs_legacy(_vp(params.h), _p(tag), _p(f), _p(cp), C.c_size_t(f.shape[0])), params, 16, 0)

    def root(self):
        out = np.zeros(4, dtype=np.uint64)
        lib().orc_tree_root(_vp(self.h), _p(out))
        return out

    def nlevels(self):
        return int(lib().orc_tree_nlevels(_vp(self.h)))

    def level(self, l):
        n = int(lib().orc_tree_level_len(_vp(self.h), int(l)))
        out = np.empty((n, 4), dtype=np.uint64)
        lib().orc_tree_level_copy(_vp(self.h), int(l), _p(out))
        return out

    def open_many(self, indices):
        idx = np.ascontiguousarray(indices, dtype=np.uint64)
        return MerkleProof(lib().orc_tree_open(_vp(self.h), _p(idx), C.c_size_t(idx.size)))

    def __del__(self):
        if self.owned:
            try:
                lib().orc_tree_free(_vp(self.h))
            except Exception:
                pass


def verify_many_ds(root, indices, values, proof, tree_label, params, arity):
    idx = np.ascontiguousarray(indices, dtype=np.uint64)
    vals = fr_array(values).reshape(-1, 4)
    if idx.size != vals.shape[0]:
        return False
    pidx = np.ascontiguousarray(proof.indices if len(proof.indices) else [0], dtype=np.uint64)
    nsib = (C.c_size_t * max(proof.nlev, 1))(*[s.shape[0] for s in proof.siblings])
    ngrp = (C.c_size_t * max(proof.nlev, 1))(*[g.shape[0] for g in proof.group_sizes])
    sib = np.concatenate([s.reshape(-1, 4) for s in proof.siblings] + [np.zeros((1, 4), dtype=np.uint64)])
    grp = np.concatenate([g for g in proof.group_sizes] + [np.zeros(1, dtype=np.uint8)])
    sib, grp = np.ascontiguousarray(sib), np.ascontiguousarray(grp)
    rt = fr_array(root).reshape(4)
    return bool(lib().orc_verify_many_ds(_p(rt), _p(idx), _p(vals), C.c_size_t(idx.size), _p(pidx), C.c_size_t(len(proof.indices)),
                                         proof.nlev, nsib, _p(sib), ngrp, grp.ctypes.data_as(u8p), int(arity), C.c_uint64(tree_label), _vp(params.h)))


def verify_many_legacy(root, indices, values, proof, ds_tag, params=None):
    """merkle/src/lib.rs:487-584 verify_many (legacy sponge in the parents; arity = proof.arity)"""
    params = params or Params.commitment_default()
    idx = np.ascontiguousarray(indices, dtype=np.uint64)
    vals = fr_array(values).reshape(-1, 4)
    if idx.size != vals.shape[0]:
        return False
    if len(proof.siblings) != len(proof.group_sizes):
        return False
    pidx = np.ascontiguousarray(proof.indices if len(proof.indices) else [0], dtype=np.uint64)
    nsib = (C.c_size_t * max(proof.nlev, 1))(*[s.shape[0] for s in proof.siblings])
    ngrp = (C.c_size_t * max(proof.nlev, 1))(*[g.shape[0] for g in proof.group_sizes])
    sib = np.ascontiguousarray(np.concatenate([s.reshape(-1, 4) for s in proof.siblings] + [np.zeros((1, 4), dtype=np.uint64)]))
    grp = np.ascontiguousarray(np.concatenate([g for g in proof.group_sizes] + [np.zeros(1, dtype=np.uint8)]))
    rt, tag = fr_array(root).reshape(4), fr_array(ds_tag).reshape(4)
    return bool(lib().orc_verify_many_legacy(_p(rt), _p(idx), _p(vals), C.c_size_t(idx.size), _p(pidx), C.c_size_t(len(proof.indices)),
                                             proof.nlev, nsib, _p(sib), ngrp, grp.ctypes.data_as(u8p), int(getattr(proof, "arity", 16) or 16), _p(tag), _vp(params.h)))


def verify_pairs_legacy(root, indices, pairs, proof, ds_tag, params=None):
    """merkle/src/lib.rs:704-721: leaf digests hash_with_ds([f, cp], ds_tag), then verify_many"""
    params = params or Params.commitment_default()
    pr = fr_array(pairs).reshape(-1, 2, 4)
    if len(indices) != pr.shape[0] or len(indices) == 0:
        return False
    leaves = np.stack([hash_with_ds_legacy(params, pr[i], ds_tag) for i in range(pr.shape[0])])
    return verify_many_legacy(root, indices, leaves, proof, ds_tag, params)


def commitment_commit(leaves, ds_tag, params=None):
    """MerkleCommitment::commit (commitment/src/lib.rs:85-90): arity 16, tree_label = ds_tag"""
    params = params or Params.commitment_default()
    return MerkleTree.new(leaves, 16, ds_tag, params)


# --------------------------------------------------------------------- fri
def fri_sample_z_ell(seed_z, level, n):
    out = np.zeros(4, dtype=np.uint64)
    lib().orc_fri_sample_z_ell(C.c_uint64(seed_z), C.c_uint64(level), C.c_uint64(n), _p(out))
    return out


def fri_fold_layer(f, z, m):
    f = fr_array(f).reshape(-1, 4)
    z = fr_array(z).reshape(4)
    out = np.empty((f.shape[0] // m, 4), dtype=np.uint64)
    rc = lib().orc_fri_fold_layer(_p(f), C.c_size_t(f.shape[0]), _p(z), C.c_size_t(m), _p(out))
    if rc:
        raise ValueError("layer size must be divisible by m")
    return out


def compute_s_layer(f, z, m):
    f = fr_array(f).reshape(-1, 4)
    z = fr_array(z).reshape(4)
    out = np.empty_like(f)
    rc = lib().orc_compute_s_layer(_p(f), C.c_size_t(f.shape[0]), _p(z), C.c_size_t(m), _p(out))
    if rc:
        raise ValueError("layer size must be divisible by m")
    return out


def pick_arity_for_layer(n, m):
    return int(lib().orc_pick_arity_for_layer(C.c_size_t(n), C.c_size_t(m)))


class FriProverState:
    """fri_build_transcript (deep_ali/src/fri.rs:231-312)"""

    def __init__(self, f0, schedule, seed_z):
        f0 = fr_array(f0).reshape(-1, 4)
        sch = np.ascontiguousarray(schedule, dtype=np.uint64)
        self.L = len(schedule)
        self.h = lib().orc_fri_build(_p(f0), C.c_size_t(f0.shape[0]), _p(sch), self.L, C.c_uint64(seed_z))

    def layer_n(self, l):
        return int(lib().orc_fri_layer_n(_vp(self.h), l))

    def layer_arity(self, l):
        return int(lib().orc_fri_layer_arity(_vp(self.h), l))

    def layer_hashed(self, l):
        return bool(lib().orc_fri_layer_hashed(_vp(self.h), l))

    def f(self, l):
        out = np.empty((self.layer_n(l), 4), dtype=np.uint64)
        lib().orc_fri_layer_f(_vp(self.h), l, _p(out))
        return out

    def s(self, l):
        out = np.empty((self.layer_n(l), 4), dtype=np.uint64)
        lib().orc_fri_layer_s(_vp(self.h), l, _p(out))
        return out

    def z(self, l):
        out = np.zeros(4, dtype=np.uint64)
        lib().orc_fri_z(_vp(self.h), l, _p(out))
        return out

    def root(self, l):
        out = np.zeros(4, dtype=np.uint64)
        lib().orc_fri_layer_root(_vp(self.h), l, _p(out))
        return out

    def roots(self):
        return np.stack([self.root(l) for l in range(self.L + 1)])

    def tree(self, l):
        return MerkleTree(lib().orc_fri_layer_tree(_vp(self.h), l), None, self.layer_arity(l), l, owned=False)

    def prove_size(self, r):
        """(proof_bytes, roots_seed, child index lists, parent index lists)"""
        rs = np.zeros(4, dtype=np.uint64)
        L = max(self.L, 1)
        ci = np.zeros((L, r), dtype=np.uint64)
        pi = np.zeros((L, r), dtype=np.uint64)
        cc = (C.c_size_t * L)()
        pc = (C.c_size_t * L)()
        sz = int(lib().orc_fri_prove_size(_vp(self.h), C.c_size_t(r), _p(rs), _p(ci), cc, _p(pi), pc))
        child = [[int(v) for v in ci[l, :cc[l]]] for l in range(self.L)]
        parent = [[int(v) for v in pi[l, :pc[l]]] for l in range(self.L)]
        return sz, rs, child, parent

    def __del__(self):
        try:
            lib().orc_fri_free(_vp(self.h))
        except Exception:
            pass


def deep_ali_merge_evals(a, s, e, t, omega, z, want_c_star=True):
    a, s, e, t = (fr_array(x).reshape(-1, 4) for x in (a, s, e, t))
    omega, z = fr_array(omega).reshape(4), fr_array(z).reshape(4)
    f0 = np.empty_like(a)
    cs = np.zeros(4, dtype=np.uint64)
    lib().orc_deep_ali_merge_evals(_p(a), _p(s), _p(e), _p(t), C.c_size_t(a.shape[0]), _p(omega), _p(z), _p(f0), _p(cs) if want_c_star else None)
    return f0, cs


def deep_ali_build_f0(a, s, e, t):
    a, s, e, t = (fr_array(x).reshape(-1, 4) for x in (a, s, e, t))
    f0 = np.empty_like(a)
    z = np.zeros(4, dtype=np.uint64)
    lib().orc_deep_ali_build_f0(_p(a), _p(s), _p(e), _p(t), C.c_size_t(a.shape[0]), _p(f0), _p(z))
    return f0, z


def deep_fri_prove_size(a, s, e, t, schedule, r, seed_z):
    a, s, e, t = (fr_array(x).reshape(-1, 4) for x in (a, s, e, t))
    sch = np.ascontiguousarray(schedule, dtype=np.uint64)
    roots = np.zeros((len(schedule) + 1, 4), dtype=np.uint64)
    sz = int(lib().orc_deep_fri_prove_size(_p(a), _p(s), _p(e), _p(t), C.c_size_t(a.shape[0]), _p(sch), len(schedule), C.c_size_t(r), C.c_uint64(seed_z), _p(roots)))
    return sz, roots


# --------------------------------------------------------------------- ntt
def ntt(a, inverse=False):
    a = fr_array(a).reshape(-1, 4).copy()
    log_n = a.shape[0].bit_length() - 1
    assert 1 << log_n == a.shape[0]
    lib().orc_ntt(_p(a), log_n, 1 if inverse else 0)
    return a


def dft_naive(coeffs):
    c = fr_array(coeffs).reshape(-1, 4)
    log_n = c.shape[0].bit_length() - 1
    out = np.empty_like(c)
    lib().orc_dft_naive(_p(c), log_n, _p(out))
    return out


def horner(coeffs, x):
    """P(x) for P = sum coeffs[j] X^j (Horner; the definition the NTT is pinned on)"""
    c = fr_array(coeffs).reshape(-1, 4)
    x = fr_array(x).reshape(4)
    out = np.zeros(4, dtype=np.uint64)
    lib().orc_horner(_p(c), C.c_size_t(c.shape[0]), _p(x), _p(out))
    return out


def fr_pow(a, e):
    a = fr_array(a).reshape(4)
    out = np.zeros(4, dtype=np.uint64)
    lib().orc_fr_pow(_p(a), C.c_uint64(e), _p(out))
    return out


def lde(col, log_blowup, shift):
    col = fr_array(col).reshape(-1, 4)
    log_n = col.shape[0].bit_length() - 1
    out = np.empty((col.shape[0] << log_blowup, 4), dtype=np.uint64)
    sh = fr_array(shift).reshape(4)
    lib().orc_lde(_p(col), log_n, int(log_blowup), _p(sh), _p(out))
    return out


# ---------------------------------------------------------------------- mf
def sumcheck_round_coeffs(layer):
    layer = fr_array(layer).reshape(-1, 4)
    c0, c1 = np.zeros(4, dtype=np.uint64), np.zeros(4, dtype=np.uint64)
    lib().orc_sumcheck_round_coeffs(_p(layer), C.c_size_t(layer.shape[0]), _p(c0), _p(c1))
    return c0, c1


def mf_round_challenge_from_root(i, prev_root, params=None):
    out = np.zeros(4, dtype=np.uint64)
    rt = fr_array(prev_root).reshape(4)
    lib().orc_mf_round_challenge_from_root(C.c_uint64(i), _p(rt), _vp(params.h) if params else None, _p(out))
    return out


def mf_fold(cur, r):
    cur = fr_array(cur).reshape(-1, 4)
    r = fr_array(r).reshape(4)
    out = np.empty((cur.shape[0] // 2, 4), dtype=np.uint64)
    lib().orc_mf_fold(_p(cur), C.c_size_t(cur.shape[0]), _p(r), _p(out))
    return out


def mf_commit_phase(table, ds_tag):
    table = fr_array(table).reshape(-1, 4)
    k = table.shape[0].bit_length() - 1
    roots = np.zeros((k + 1, 4), dtype=np.uint64)
    rs = np.zeros((max(k, 1), 4), dtype=np.uint64)
    fin = np.zeros(4, dtype=np.uint64)
    lib().orc_mf_commit_phase(C.c_uint64(ds_tag), _p(table), C.c_size_t(table.shape[0]), _p(roots), _p(rs), _p(fin))
    return roots, rs[:k], fin
