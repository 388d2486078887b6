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
 * Launch geometry: grid = ceil(n_chains / chains_per_block) CTAs of <block>
 * threads, dynamic smem = smem_fixed + flen[smem_field] * bytes_per_elem.
 *
 * State layout in HBM: field f of per-chain length L is [L][n_chains]
 * (chain-minor) — structure-of-arrays, coalesced across chains.
 */
#ifndef STK_ABI_H
#define STK_ABI_H

#define STK_MAX_FIELDS 24
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
};

#endif /* STK_ABI_H */
