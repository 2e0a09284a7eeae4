"""Loads libsmw.so and declares the C-ABI signatures (include/smw.h)."""
import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.environ.get("SMW_LIB") or os.path.join(_HERE, "libsmw.so")   # SMW_LIB: a tuning build of the same library
HEADER = os.path.join(_HERE, "..", "include", "smw.h")

OK = 0
HOST, DEVICE = 0, 1
ERR = {1: "SMW_ERR_CUDA", 2: "SMW_ERR_EMPTY", 3: "SMW_ERR_ARITY_WIDTH", 4: "SMW_ERR_LENGTH_MISMATCH", 5: "SMW_ERR_NOT_DIVISIBLE",
       6: "SMW_ERR_UNSUPPORTED_WIDTH", 7: "SMW_ERR_UNSUPPORTED_ARITY", 8: "SMW_ERR_INVALID_ARG", 9: "SMW_ERR_PARAMS_SINGULAR",
       10: "SMW_ERR_NOMEM", 11: "SMW_ERR_EMPTY_INDICES", 12: "SMW_ERR_Z_IN_DOMAIN", 13: "SMW_ERR_COMM"}

_lib = None


def declared_symbols():
    """every function name declared in include/smw.h"""
    txt = open(HEADER).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(smw_[a-z0-9_]+)\s*\(", txt)))


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError("stark_mlwe_b200: %s is missing — build it with `python -m stark_mlwe_b200.build` "
                          "(there is no CPU fallback)" % SO_PATH)
    L = C.CDLL(SO_PATH)
    vp, sz, u64, i32 = C.c_void_p, C.c_size_t, C.c_uint64, C.c_int32
    for name in declared_symbols():
        getattr(L, name)   # AttributeError if the library does not export a declared symbol
    L.smw_last_error.restype = C.c_char_p
    L.smw_ctx_launch_count.restype = u64
    L.smw_ctx_collective_count.restype = u64
    L.smw_ctx_stream.restype = vp
    for n in ("smw_tree_level_len", "smw_fri_shard_tail_len", "smw_proof_num_indices", "smw_proof_level_siblings", "smw_proof_level_groups"):
        getattr(L, n).restype = sz
    L.smw_fri_layer_tree.restype = vp
    for n in ("smw_ctx_destroy", "smw_params_free", "smw_tree_free", "smw_transcript_free", "smw_fri_free", "smw_fri_shard_free", "smw_proof_free", "smw_comm_free"):
        getattr(L, n).restype = None
    # argument types: every pointer is passed as c_void_p, sizes as size_t, labels/seeds as uint64
    A = {
        "smw_merkle_new_legacy": [vp, vp, vp, vp, sz, i32, vp], "smw_merkle_new_pairs_legacy": [vp, vp, vp, vp, vp, sz, sz, i32, vp],
        "smw_verify_many_legacy": [vp, vp, i32, vp, vp, vp, vp, sz, vp, sz, i32, vp, vp, vp, vp, vp],
        "smw_verify_pairs_legacy": [vp, vp, i32, vp, vp, vp, vp, sz, vp, sz, i32, vp, vp, vp, vp, vp],
        "smw_comm_unique_id": [vp], "smw_comm_init": [vp, vp, i32, i32, vp], "smw_comm_rank": [vp], "smw_comm_world": [vp], "smw_comm_free": [vp],
        "smw_comm_all_gather": [vp, vp, vp, sz], "smw_ctx_collective_count": [vp], "smw_ctx_set_tc_min_nodes": [vp, u64], "smw_diag_tc_node_counts": [vp, vp, vp], "smw_sumcheck_round_coeffs_noalloc": [vp, vp, sz, vp],
        "smw_fri_roots_sharded": [vp, vp, vp, sz, i32, vp, sz, vp, i32, u64, i32, vp],
        "smw_ntt_cyclic_to_chunks": [vp, vp, vp, vp, i32],
        "smw_mf_commit_phase_sharded": [vp, vp, u64, vp, sz, i32, vp, vp, vp],
        "smw_ctx_create": [i32, vp], "smw_ctx_create_on_stream": [i32, vp, vp], "smw_ctx_destroy": [vp], "smw_ctx_sync": [vp], "smw_ctx_release_cached": [vp],
        "smw_last_error": [vp], "smw_ctx_launch_count": [vp], "smw_ctx_stream": [vp],
        "smw_ctx_profile_enable": [vp, i32], "smw_ctx_profile_read": [vp, vp, vp, i32], "smw_diag_mac_rate": [vp, vp, vp], "smw_diag_fp64_rate": [vp, vp, vp], "smw_diag_pipe_rate": [vp, i32, vp], "smw_diag_perm_fp64_ops": [vp, vp], "smw_diag_f52_dot": [vp, vp, vp, sz, i32, vp, i32], "smw_index_from_seed": [vp, vp, sz, u64, vp, i32], "smw_ali_sample_z_beta": [vp, vp, vp, u64, vp, vp], "smw_scale_powers": [vp, vp, sz, vp, i32], "smw_dft_strided": [vp, vp, i32, sz, sz, i32],
        "smw_fri_build_transcript_from": [vp, vp, sz, vp, i32, u64, u64, i32, vp],
        "smw_fri_shard_build": [vp, vp, sz, sz, sz, vp, i32, u64, i32, vp], "smw_fri_shard_num_local_layers": [vp],
        "smw_fri_shard_cap_info": [vp, i32, vp, vp, vp], "smw_fri_shard_cap": [vp, i32, vp, i32], "smw_fri_shard_tail_len": [vp],
        "smw_fri_shard_tail": [vp, vp, i32], "smw_fri_shard_free": [vp],
        "smw_merkle_root_from_level": [vp, vp, i32, u64, i32, vp, sz, i32, vp], "smw_diag_perm_macs": [vp, vp],
        "smw_fr_from_canonical": [vp, vp, vp, sz, i32], "smw_fr_to_canonical": [vp, vp, vp, sz, i32], "smw_root_of_unity": [vp, i32, vp],
        "smw_params_from_seed": [vp, vp, sz, i32, i32, vp], "smw_params_for_width": [vp, i32, vp], "smw_params_for_arity": [vp, i32, vp],
        "smw_params_commitment_default": [vp, vp], "smw_params_transcript_default": [vp, vp],
        "smw_params_from_tables": [vp, i32, i32, vp, vp, vp, i32, vp], "smw_params_export": [vp, vp, vp, vp],
        "smw_params_width": [vp], "smw_params_partial_rounds": [vp], "smw_params_free": [vp],
        "smw_poseidon_permute": [vp, vp, vp, sz, i32], "smw_poseidon_permute_dense": [vp, vp, vp, sz, i32],
        "smw_hash_with_ds_dynamic": [vp, vp, vp, sz, vp, sz, vp, i32], "smw_hash_with_ds": [vp, vp, vp, sz, vp, vp, i32],
        "smw_merkle_new": [vp, vp, i32, u64, vp, sz, i32, vp], "smw_merkle_new_pairs": [vp, vp, i32, u64, vp, vp, sz, sz, i32, vp],
        "smw_tree_root": [vp, vp], "smw_tree_num_levels": [vp], "smw_tree_level_len": [vp, i32], "smw_tree_level": [vp, i32, vp, i32],
        "smw_tree_arity": [vp], "smw_tree_free": [vp],
        "smw_merkle_commit_columns": [vp, vp, i32, u64, u64, vp, sz, i32, i32, vp],
        "smw_commitment_commit": [vp, u64, vp, vp, sz, i32, vp],
        "smw_ntt": [vp, vp, i32, i32, i32, i32], "smw_lde": [vp, vp, i32, i32, vp, vp, i32, i32],
        "smw_lde_commit": [vp, vp, i32, i32, vp, i32, u64, i32, vp],
        "smw_transcript_new": [vp, vp, sz, vp, vp], "smw_transcript_absorb_bytes": [vp, vp, sz], "smw_transcript_absorb_field": [vp, vp],
        "smw_transcript_absorb_fields": [vp, vp, sz, i32], "smw_transcript_challenge": [vp, vp, sz, vp], "smw_transcript_state": [vp, vp, vp],
        "smw_transcript_free": [vp], "smw_tr_hash_fields_tagged": [vp, vp, sz, vp, sz, i32, vp], "smw_hash_leaf_pair": [vp, vp, vp, sz, vp, i32],
        "smw_fri_sample_z_ell": [vp, u64, u64, u64, vp], "smw_fri_fold_layer": [vp, vp, sz, vp, sz, vp, i32],
        "smw_compute_s_layer": [vp, vp, sz, vp, sz, vp, i32], "smw_pick_arity_for_layer": [sz, sz],
        "smw_fri_build_transcript": [vp, vp, sz, vp, i32, u64, i32, vp], "smw_fri_num_layers": [vp],
        "smw_fri_layer_info": [vp, i32, vp, vp, vp, vp], "smw_fri_layer_f": [vp, i32, vp, i32], "smw_fri_layer_s": [vp, i32, vp, i32],
        "smw_fri_roots": [vp, vp], "smw_fri_z": [vp, vp], "smw_fri_layer_tree": [vp, i32], "smw_fri_free": [vp],
        "smw_sumcheck_round_coeffs": [vp, vp, sz, vp, vp, i32], "smw_mf_round_challenge_from_root": [vp, u64, vp, vp, vp],
        "smw_mf_fold": [vp, vp, sz, vp, vp, i32], "smw_mf_commit_phase": [vp, u64, vp, sz, i32, vp, vp, vp],
        "smw_tree_open_many": [vp, vp, sz, vp], "smw_proof_num_levels": [vp], "smw_proof_num_indices": [vp], "smw_proof_indices": [vp, vp],
        "smw_proof_level_siblings": [vp, i32], "smw_proof_level_groups": [vp, i32], "smw_proof_level": [vp, i32, vp, vp], "smw_proof_free": [vp],
        "smw_deep_ali_merge_evals": [vp, vp, vp, vp, vp, sz, vp, vp, vp, i32],
        "smw_tr_hash_fields_tagged_many": [vp, i32, vp, vp, vp, vp, i32, vp],
        "smw_tr_hash_fields_tagged_batch": [vp, vp, sz, vp, i32, sz, i32, vp],
        "smw_verify_many_ds": [vp, vp, i32, u64, vp, vp, vp, sz, vp, sz, i32, vp, vp, vp, vp, vp],
        "smw_verify_pairs_ds": [vp, vp, i32, u64, vp, vp, vp, sz, vp, sz, i32, vp, vp, vp, vp, vp],
    }
    for n, a in A.items():
        getattr(L, n).argtypes = a
    _lib = L
    return L
