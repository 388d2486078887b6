/* stk_abi.h — the contract between libstk_cuda's host runtime and a compiled
 * model module (a cubin: either a hand-tuned builtin or NVRTC output of the
 * CUDA text that stkc / Language.CUDA generates).  Plain C, no includes, so
 * that NVRTC can compile it without a host toolchain.
 *
 * A module exports
 *   __device__ const char stk_desc[]   the descriptor text (grammar below)
 *   extern "C" __global__ kernels      each taking ONE argument: StkArgs by
 *                                      value (__grid_constant__)
 *
 * Descriptor grammar (one record per line, blank-separated):
 *   stk-module 1
 *   name   <model name>
 *   nparam <P>                       unconstrained parameter count (0 if n/a)
 *   field  <name> <state|data> <f64|i32> <len> [sorted]   len 0 = sized by set_data; @f = as field f;
 *                                    sorted = the runtime uploads an ascending copy of the data
 *   param  <name> <default>          scalar tunable -> StkArgs.par[i]
 *   record <field> ...               fields (as doubles) that make up a draw
 *   kernel <name> <symbol> <block> <chains_per_block> <smem_fixed> <smem_field|-> <smem_bytes_per_elem>
 *          [pre <symbol> <block> <doubles_per_chain_transition>]
 *          [phase <symbol> <block> <chains_per_block> <items_field> <items_per_block> <smem_bytes>]
 *          transition kernels, kernel_id = order of appearance.  `pre` names a pre-pass kernel that the runtime
 *          launches first, one thread per (chain, transition) of the launch, to fill StkArgs.scratch
 *          [transition][k][chain] (k < doubles_per_chain_transition) — the state-independent part of a transition
 *          (its random draws) computed at full occupancy instead of inside the sequential chain loop.
 *          `phase` names a kernel that runs before the main kernel in EVERY transition (the launch is then split
 *          into single transitions) on a 2-D grid: x = ceil(n_chains / chains_per_block) chain groups,
 *          y = ceil(flen[items_field] / items_per_block) tiles of the per-item work of one transition (e.g. the
 *          assignment step of a Gibbs sweep over the observations); the main kernel then finishes the transition.
 *   entry  <logp|logp_grad|simulate> <symbol> <block> <chains_per_block> <smem_fixed> <smem_field|-> <smem_bytes_per_elem>
 *   plan   <name> <kernel|logp|logp_grad|simulate|prepare> <loop-param|->
 *   stage  <plan> <init|pre|body|post> <symbol> <block> <gx> <gy> <smem> [coop]
 *          A plan is a transition kernel (or an entry) made of several launches — what a model needs when one
 *          transition is a sequence of grid-wide steps (a matrix product over all chains, then an elementwise
 *          update, ...).  Per stk_step call the runtime runs the `init` stages once if the chains are dirty
 *          (data or state uploaded since the last launch); then per transition: the `pre` stages, par[loop-param]
 *          times the `body` stages (StkArgs.iter = 0.., StkArgs.n_iter = the count), the `post` stages.  An entry
 *          plan runs its stages once, in order.  A `prepare` plan runs before any other launch whenever a data
 *          field changed (per-data-set precomputation into `work` fields); if the module has an i32 work field
 *          named `info`, a non-zero info[0] after prepare is reported as an error.
 *          gx / gy: <int> | sms | chains/<k> | @<field>/<k> (ceiling division; `sms` = SM count of the device);
 *          `coop` launches cooperatively (grid-wide sync allowed, grid must be co-resident).
 * Field kinds: `state` (per chain, [len][n_chains]), `data` (uploaded, shared by all chains), `work` (runtime-
 * allocated scratch shared by all chains, len = total elements).  Types f64 | i32 | f32 | bf16 | f16.  <len> is a product
 * of terms <int> | @<field> | @<field>^<k> (length of the field rounded up to a multiple of k) | chains | chains^<k>.
 * `nparam` accepts the same expressions (models whose parameter count follows the data).
 * Launch geometry: grid = ceil(n_chains / chains_per_block) CTAs of <block>
 * threads, dynamic smem = smem_fixed + flen[smem_field] * bytes_per_elem.
 *
 * State layout in HBM: field f of per-chain length L is [L][n_chains]
 * (chain-minor) — structure-of-arrays, coalesced across chains.
 */
#ifndef STK_ABI_H
#define STK_ABI_H

#define STK_MAX_FIELDS 48
#define STK_MAX_PARAMS 16

struct StkArgs {
  long long n_chains;          /* chains in this shard                        */
  long long chain0;            /* global index of the shard's first chain     */
  unsigned long long seed;     /* Philox key                                  */
  unsigned long long step0;    /* transition counter at launch start          */
  long long n_steps;           /* transitions to run in this launch           */
  void* field[STK_MAX_FIELDS]; /* device pointers, descriptor order           */
  long long flen[STK_MAX_FIELDS]; /* per-chain length (state) / total (data)  */
  double par[STK_MAX_PARAMS];  /* scalar tunables, descriptor order           */
  double* draws;               /* [draw_slots][P_rec][n_chains] or NULL       */
  long long draw_slots;        /* capacity of the ring                        */
  long long draw_thin;         /* keep every thin-th transition               */
  long long draw_base;         /* transitions recorded before this launch     */
  long long n_rec;             /* doubles per draw per chain (P_rec)          */
  double* out_lp;              /* [n_chains]   (logp / logp_grad entries)     */
  double* out_grad;            /* [P][n_chains]                               */
  unsigned long long* accept;  /* [n_chains] accepted-transition counters     */
  double* scratch;             /* pre-pass output [n_steps][k][n_chains], or NULL */
  long long iter;              /* plan stages: index of the body iteration    */
  long long n_iter;            /* plan stages: body iterations per transition (0 outside a transition plan) */
  int sm_count;                /* SMs of the device                           */
  int stage_kind;              /* 0 init, 1 pre, 2 body, 3 post, 4 entry/prepare */
};

#endif /* STK_ABI_H */
