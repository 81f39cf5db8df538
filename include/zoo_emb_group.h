/* Grouped embedding kernels (csrc/ops/embedding_group.cu): plain-C parameter blocks shared by the kernels and their callers.
 *
 * An "id column group" is one column of the id matrix together with every table that is looked up with it (NeuralCF: the
 * user column feeds the MLP-user and the MF-user table, the item column the two item tables).  All tables of a group see
 * the same ids, so the backward elects ONE representative entry per touched row for the whole group, keeps ONE winner list
 * and applies the sparse optimizer to the rows of every table of the group in the same launch; and all groups of a step
 * share each launch (blockIdx.y = group), which keeps independent latency-bound work in flight together.
 * Replaces, like csrc/ops/embedding.cu: LookupTable / LookupTableSparse inside zoo Embedding
 * (zoo/.../pipeline/api/keras/layers/Embedding.scala:82-101), the GMF merge of NeuralCF.scala:83-89 and the dense-table
 * all-reduce + optimizer the reference performs for embedding gradients (Topology.scala:1187-1206).
 */
#ifndef ZOO_EMB_GROUP_H_
#define ZOO_EMB_GROUP_H_

#define ZOO_EG_MAX_PEERS 8
#define ZOO_EG_MAX_GROUPS 4
#define ZOO_EG_MAX_TABS 2
#define ZOO_EG_MAX_SPECS 4

#ifdef __cplusplus
extern "C" {
#endif

/* One table of a column group, as seen by the rank that owns the shard. */
typedef struct ZooEgTab {
  float* w;                 /* fp32 master rows [rows_local, D] */
  void* w16;                /* optional bf16 shadow rows (read by forward gathers) */
  float* m;                 /* Adam first moment / Adagrad accumulator (null for SGD) */
  float* v;                 /* Adam second moment (null otherwise) */
  float* accum;             /* fp32 [max_entries, D], zero on entry, restored on exit */
  float* hyper;             /* device fp32 [>= 6]: lr, beta1, beta2, eps, weight_decay, step */
  const void* dout[ZOO_EG_MAX_PEERS];   /* per SOURCE rank: bf16 gradient rows of this table's window, row b at + b * ldg */
  const void* mul[ZOO_EG_MAX_PEERS];    /* per source rank: optional bf16 factor rows (GMF product rule), all null if unused */
  long long ldg, ldmul;
  int D;                    /* row width, multiple of 4 */
  int opt;                  /* 0 SGD, 1 Adam (touched rows only), 2 Adagrad */
  float grad_scale;
  int bump_step;            /* 1: hyper[5] += 1 before it is used (device-side optimizer step counter) */
} ZooEgTab;

typedef struct ZooEgGroup {
  const long long* ids;     /* this rank's id matrix [B, id_stride] (world == 1: THE entry list) */
  const void* pairs;        /* world > 1: int2 [world * cap] = (local row, entry index at the source), filled by zoo_eg_route */
  const int* counts;        /* world > 1: int32 [world] valid pairs per source */
  int* rep_of_row;          /* int32 [rows_local], INT_MAX everywhere on entry, restored on exit */
  int* dup_flag;            /* int32 [ntabs * max_entries] (one flag array per table), zero on entry, restored on exit */
  void* win_list;           /* int2 [max_entries] scratch (any contents) */
  int* win_count;           /* int32 [1] scratch */
  long long rows;           /* global (unsharded) row count */
  int id_stride, id_col, bag, E_local;   /* E_local = B * bag */
  int cap, world, rank, ntabs;
  int max_entries, pad_;    /* capacity of the scratch arrays: >= B * bag (world == 1), >= the entries this rank can own (world > 1) */
  ZooEgTab t[ZOO_EG_MAX_TABS];
} ZooEgGroup;

typedef struct ZooEgParams {
  int ngroups, pad_;
  ZooEgGroup g[ZOO_EG_MAX_GROUPS];
} ZooEgParams;

/* One lookup of a fused gather launch. */
typedef struct ZooEgSpec {
  const void* ta[ZOO_EG_MAX_PEERS];     /* shard pointers of the table (index = owner rank) */
  const void* tb[ZOO_EG_MAX_PEERS];     /* second table of a product lookup (all null otherwise) */
  void* out;                /* bf16 window start, row b at + b * ldo */
  void* save_a;             /* optional bf16 [B, D]: the two factors of a product (bag must be 1) */
  void* save_b;
  long long rows_a, rows_b, ldo;
  int id_col, id_col2, bag, mean, D, dtype;   /* dtype: 0 fp32 rows, 1 bf16 rows */
} ZooEgSpec;

typedef struct ZooEgGatherParams {
  const long long* ids;
  int id_stride, B, world, nspecs;
  ZooEgSpec s[ZOO_EG_MAX_SPECS];
} ZooEgGatherParams;

/* Source-side routing of one id column to the owners of its rows. */
typedef struct ZooEgRouteGroup {
  void* pairs[ZOO_EG_MAX_PEERS];        /* every rank's pairs buffer (symmetric heap), int2 [world * cap] */
  int* counts[ZOO_EG_MAX_PEERS];        /* every rank's counts buffer, int32 [world] */
  int* counters;                        /* local int32 [ZOO_EG_MAX_PEERS + 1], zero on entry, restored on exit ([8] = CTA ticket) */
  long long rows;
  int id_col, bag;
} ZooEgRouteGroup;

typedef struct ZooEgRouteParams {
  const long long* ids;
  int id_stride, B, world, rank, cap, ngroups;
  ZooEgRouteGroup g[ZOO_EG_MAX_GROUPS];
} ZooEgRouteParams;

/* Forward: all lookups of a step in ONE launch (blockIdx.y = spec).  Rows are read straight from their owner's shard through
 * peer pointers, as in zoo_embedding_gather. */
int zoo_eg_gather(const ZooEgGatherParams* p, void* stream);
/* Forward, world > 1: every source tells each owner which of its rows the batch touches -- (local row, entry index) pairs
 * written into the owner's `pairs[rank * cap ...]` with peer stores, and the pair count into `counts[rank]` -- so the owner's
 * backward passes run over ~B compact entries instead of scanning world * B ids.  cap >= B * bag. */
int zoo_eg_route(const ZooEgRouteParams* p, void* stream);
/* Backward, after a device barrier when world > 1 (every rank's gradient rows are staged, every rank's forward gathers are
 * done): elect representatives, sum duplicate gradient rows, apply the sparse optimizer -- three launches for all groups and
 * all their tables.  Gradient rows of remote entries are read in place over NVLink (the backward all-to-all happens inside
 * these kernels). */
int zoo_eg_update(const ZooEgParams* p, void* stream);
/* sizeof of {ZooEgTab, ZooEgGroup, ZooEgParams, ZooEgSpec, ZooEgGatherParams, ZooEgRouteParams}: lets a binding in another
 * language check its mirror of these structs. */
int zoo_eg_sizes(int* out6);

#ifdef __cplusplus
}
#endif
#endif /* ZOO_EMB_GROUP_H_ */
