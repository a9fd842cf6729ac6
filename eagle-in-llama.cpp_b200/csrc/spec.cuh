// spec.cuh — device-resident state of one speculative-decoding stream.
#pragma once
#include <stdint.h>
#include "../../include/eagle_b200.h"

struct FwdCtl { int n_prefix; int node_base; int n_tree; };   // read by the engine kernels on device

struct SpecCfg { int max_nodes, n_branch, n_draft, max_depth, top_k, fixed_accept; float p_split; float p_min; };

struct SpecState {
    // committed state
    int n_past; int root_token;
    // tree of the current round (node 0 = root)
    int n_nodes; int tokens[128]; int parent[128]; int pos[128]; int depth[128];
    alignas(16) uint32_t tree_bits[128 * 4];   // ancestor-or-self sets (read as uint4)
    // branches (the reference's drafts[s], speculative-eagle.cpp:21-33)
    int n_seq_cur; int br_drafting[16]; int br_last_node[16]; int br_row[16];
    int cur_depth; int done;
    // draft batch for the next head forward
    int gate_fwd, gate_lm;      // launch gates of the next head forward / of its lm_head + top-k (common.cuh)
    int dft_n; int dft_node_base; int dft_tokens[16]; int dft_pos[16]; int dft_src[16];
    alignas(16) uint32_t dft_bits[16 * 4];
    // accept
    int n_accept; int bonus; int last_node; int path[128];
    FwdCtl ctl_head, ctl_tgt;
};

int eb_spec_round_begin(SpecState *s, const SpecCfg &c, cudaStream_t st);
int eb_spec_gather_hidden(const SpecState *s, const float *root_feat, const float *node_feat, float *hidden, int E, int n_branch, cudaStream_t st);
int eb_spec_scatter_feat(const SpecState *s, const float *feat, float *node_feat, int E, int n_branch, cudaStream_t st);
int eb_spec_expand(SpecState *s, const SpecCfg &c, const int32_t *cand_ids, const float *cand_p, cudaStream_t st);
int eb_spec_accept(SpecState *s, const SpecCfg &c, const int32_t *best, cudaStream_t st);
int eb_spec_compact(SpecState *s, void *kc, void *vc, void *stage, int64_t layer_stride, int E, int L, int max_acc, cudaStream_t st);
int eb_spec_round_end(SpecState *s, const float *tgt_feat, float *root_feat, int E, eb_round_result *res, cudaStream_t st);
int eb_top_k_impl(const float *logits, int rows, int V, int k, int32_t *ids, float *vals, float *probs, cudaStream_t st, int ks = 1, int64_t plane = 0, void *scratch = nullptr);
size_t eb_top_k_scratch_bytes(int rows);
void *eb_top_k_default_scratch(cudaStream_t st);
int eb_argsort_impl(const float *x, int rows, int n, int descending, int32_t *ids, cudaStream_t st);
int eb_argmax_impl(const float *logits, int rows, int V, int32_t *ids, cudaStream_t st, int ks = 1, int64_t plane = 0);
