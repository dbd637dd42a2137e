/*
 * oe_abi.h — C ABI of the B200 offer-evaluation engine.
 *
 * This is the drop-in boundary for ONE path of the DC/OS SDK scheduler: matching Mesos offers against
 * pod-instance requirements.  Citations use the shorthands
 *   S/ = sdk/scheduler/src/main/java/com/mesosphere/sdk/      P/ = S/offer/evaluate/placement/
 * and point at the reference interface each entry point or table replaces.
 *
 * Conventions
 *   - extern "C", plain-old-data, struct-of-arrays.  No torch / C++ types cross this boundary.
 *   - Every input pointer is a HOST pointer owned by the caller; the engine copies what it needs into its own
 *     device memory during oe_load_*().  Outputs are written into caller-owned host buffers.
 *   - Strings never cross the boundary: the host dictionary-encodes agent ids, hostnames, zones, regions,
 *     attribute strings, pod types, task names, volume profiles and resource ids into dense int32 ids
 *     (0..n-1, -1 = absent) and pre-resolves every StringMatcher (P/StringMatcher.java:15) over every
 *     dictionary value into bitmasks (`matcher_bits`).
 *   - Return value: 0 = OK, <0 = oe_status error; oe_last_error() gives the text.  Nothing aborts or throws;
 *     a JNI shim maps non-zero to IOException so that S/scheduler/plan/PlanScheduler.java:99-104 handles it
 *     like any other evaluation failure ("no recommendations for this step").
 *   - One engine handle per service, driven from one thread (the reference evaluates on the single
 *     `offerExecutor` thread, S/framework/OfferProcessor.java:57) — no internal locking.
 *   - 2-D columns are laid out slot-major: col[slot][row] = col[slot * n_rows + row], so that consecutive rows
 *     of one slot are consecutive in memory (coalesced on the device).
 */
#ifndef OE_ABI_H_
#define OE_ABI_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define OE_ABI_VERSION 1

typedef enum {
  OE_OK = 0,
  OE_ERR_INVALID = -1,   /* malformed table / argument out of range */
  OE_ERR_STATE = -2,     /* call order violated (e.g. evaluate before load) */
  OE_ERR_CUDA = -3,      /* a CUDA runtime call failed */
  OE_ERR_NO_DEVICE = -4, /* no usable sm_100 device: there is no CPU fallback */
  OE_ERR_LIMIT = -5      /* a documented fixed limit was exceeded */
} oe_status;

/* ------------------------------------------------------------------------------------------------------------------
 * Limits (engine rejects tables beyond them with OE_ERR_LIMIT)
 * ---------------------------------------------------------------------------------------------------------------- */
#define OE_MAX_ROLE_SLOTS 4     /* distinct reservable roles seen in one cycle ("*" is slot 0)            */
#define OE_MAX_SCALAR_KINDS 8   /* cpus, mem, disk, gpus + custom scalar resource names                   */
#define OE_MAX_PORT_RANGES 32   /* merged `ports` ranges per (offer, role slot)                           */
#define OE_MAX_MOUNT_SLOTS 8    /* unreserved MOUNT disks per offer                                       */
#define OE_MAX_ATTR_SLOTS 16    /* attributes per offer / per task                                        */
#define OE_MAX_RR_COLS 8        /* distinct RoundRobinByAttribute attribute names                         */
#define OE_MAX_EXEC_SLOTS 8     /* executor ids listed in one offer                                       */
#define OE_MAX_STAGE_OPS 64     /* stage ops per requirement class                                        */
#define OE_MAX_RULE_NODES 64    /* postfix rule nodes per requirement class                               */
#define OE_MAX_CLASSES 4096     /* requirement classes per cycle                                          */
#define OE_MAX_DYN_PORTS 8      /* dynamic host ports per class (reported in oe_decision)                 */
#define OE_MAX_MOUNT_OPS 4      /* MOUNT volumes per class (reported in oe_decision)                      */
#define OE_MAX_PROFILES 64      /* volume profile dictionary (profile sets are u64 masks)                 */

#define OE_PORT_EMPTY 0x0000FFFFu /* packed range with begin(65535) > end(0): unused range slot           */
#define OE_PORT_PACK(b, e) ((uint32_t)(b) | ((uint32_t)(e) << 16))
#define OE_NO_INDEX INT32_MIN     /* task row without a parsable index label                               */

/* ------------------------------------------------------------------------------------------------------------------
 * Engine set-up
 * ---------------------------------------------------------------------------------------------------------------- */
typedef struct oe_engine oe_engine;

typedef struct {
  int32_t device;       /* CUDA device ordinal                                                             */
  int32_t rank;         /* offer-shard index of this engine (0 when not sharded)                           */
  int32_t world;        /* number of offer shards (1 when not sharded)                                     */
  uint32_t flags;       /* OE_CFG_*                                                                        */
  void* stream;         /* cudaStream_t to launch on; NULL = engine creates its own non-blocking stream    */
} oe_config;

#define OE_CFG_DEFAULT 0u
/* oe_evaluate_cycle also writes the full offers x requirements bit matrix into OE_BUF_MATRIX (K1x) */
#define OE_CFG_MATERIALIZE_MATRIX 1u
/* oe_load_* only ENQUEUE their host→device copies and return; the caller keeps every table buffer alive and unchanged until
 * the next evaluation call on the handle returns (that call reports a malformed id column, which a synchronous load would
 * have reported itself).  With pinned host buffers the upload of the offer table then overlaps the host-side part of the
 * other loads AND the evaluation: the table travels in row chunks, every chunk is checked when it has landed, and the next
 * oe_evaluate_cycle / oe_feasibility / oe_class_feasibility evaluates the landed chunks while the later ones are still on
 * the bus (oe_evaluate_cycle also commits over them: first fit takes the lowest free index, so what a requirement gets among
 * the landed offers does not depend on the offers behind them).  Load the offers before the requirements to use it.
 * With this flag the loads also do nothing per row on the host: ids are checked, and requirements that re-deploy a running
 * pod found, on the device; what oe_apply_launches / oe_requirement_rows need of the tables on the host is fetched when they
 * are called (from the caller's buffers while the promise above holds, from the device afterwards — the engine keeps no
 * pointer into a caller's buffer once that evaluation call has returned).
 * Without this flag a load returns only when the caller may reuse its buffers.                                           */
#define OE_CFG_ASYNC_LOADS 2u

int32_t oe_abi_version(void);
/* Replaces the construction of S/offer/evaluate/OfferEvaluator.java:85-111 (one per service,
 * S/scheduler/DefaultScheduler.java:198-210).  Fails with OE_ERR_NO_DEVICE when no CUDA device is usable. */
int32_t oe_create(const oe_config* cfg, oe_engine** out);
void oe_destroy(oe_engine* h);
/* Text of the last error on this handle (or of the last failed oe_create when h == NULL). */
const char* oe_last_error(const oe_engine* h);

/* ------------------------------------------------------------------------------------------------------------------
 * Dictionary sizes: id spaces shared by the tables of one cycle
 * ---------------------------------------------------------------------------------------------------------------- */
typedef struct {
  int32_t n_agents;     /* Offer.slave_id / TaskInfo.slave_id values                                       */
  int32_t n_hostnames;  /* Offer.hostname / task label offer_hostname                                      */
  int32_t n_zones;      /* fault-domain zone names                                                         */
  int32_t n_regions;    /* fault-domain region names                                                       */
  int32_t n_attrs;      /* attribute strings "name:value" (S/offer/taskdata/AttributeStringUtils.java:81)  */
  int32_t n_rr_values;  /* attribute VALUE strings used as RoundRobinByAttribute keys                      */
  int32_t n_pod_types;  /* pod type names (task label task_type)                                           */
  int32_t n_task_names; /* task names (state-store key, S/state/StateStore.java:340)                       */
  int32_t n_filters;    /* distinct task-filter StringMatchers referenced by counter sets                  */
  int32_t n_exec_ids;   /* executor ids                                                                    */
  int32_t n_roles;      /* role names (allocation roles / pod roles)                                       */
  int32_t n_res_ids;    /* resource-id labels of dynamically reserved resources                            */
} oe_dict_sizes;

/* ------------------------------------------------------------------------------------------------------------------
 * Offer table — N rows in LIST ORDER (first-fit index order is semantic, S/offer/evaluate/OfferEvaluator.java:137).
 * One row is what `new MesosResourcePool(offer, role)` (S/offer/MesosResourcePool.java:53-59) plus the placement
 * getters read from one Protos.Offer.
 * ---------------------------------------------------------------------------------------------------------------- */
typedef struct {
  int64_t n_offers;        /* rows in this table (the local shard when world > 1)                            */
  int64_t row_offset;      /* global list index of row 0 (0 when not sharded)                                */
  int64_t n_offers_global; /* rows over all shards (== n_offers when not sharded)                            */
  int64_t shard_rows;      /* uniform shard capacity when world > 1: row_offset == rank * shard_rows, a multiple
                              of 1024; every shard but the last holds exactly shard_rows rows.  0 = not sharded */
  int32_t n_role_slots;    /* S                                                                              */
  int32_t n_scalar_kinds;  /* K                                                                              */
  int32_t n_port_ranges;   /* R                                                                              */
  int32_t n_mount_slots;   /* D                                                                              */
  int32_t n_attr_slots;    /* A                                                                              */
  int32_t n_rr_cols;       /* Q                                                                              */
  int32_t n_exec_slots;    /* E                                                                              */
  int32_t reserved0;

  /* reservableMergedPoolByRole (MesosResourcePool.java:432-455): fp64 sum, in offer-resource order (:133-147), of
   * the non-atomic resources without resource-id, per role slot and scalar kind.  [S][K][N]                    */
  const double* scalar;
  /* bit (s*K + k) set iff the pool of slot s holds an entry for kind k (null entry ⇒ sufficientValue false,
   * MesosResourcePool.java:61-66).  [N]                                                                         */
  const uint32_t* scalar_present;
  /* merged `ports` ranges of slot s, sorted, disjoint, adjacency-coalesced (antlr IntervalSet via
   * S/offer/RangeUtils.java:23-47), packed OE_PORT_PACK(begin,end); unused = OE_PORT_EMPTY.  [S][R][N]         */
  const uint32_t* port_range;
  /* unreservedAtomicPool["disk"] (MesosResourcePool.java:99-117): MOUNT disks without resource-id in offer order;
   * absent slot = -inf.  [D][N]                                                                                  */
  const double* mount_size;
  const int32_t* mount_profile; /* profile id or -1 (MesosResourcePool.java:289-292).  [D][N]                  */
  /* allocation_info.role of the offer's resources, -1 = not set (consumableResource, MesosResourcePool.java:87-97).
   * Precondition: all resources of one offer share one allocation role (Mesos allocates an offer to one role).  */
  const int32_t* alloc_role;    /* [N] */
  const int32_t* agent_id;      /* Offer.slave_id  [N] */
  const int32_t* hostname_id;   /* Offer.hostname  [N] */
  const int32_t* zone_id;       /* Offer.domain.fault_domain.zone.name, -1 = offer has no fault domain  [N]     */
  const int32_t* region_id;     /* ... region.name, -1 = no fault domain  [N]                                   */
  const int32_t* attr_id;       /* attribute strings in offer order (duplicates kept), -1 pad.  [A][N]          */
  const int32_t* rr_key;        /* value id of the first attribute whose name equalsIgnoreCase column q's name
                                   (P/RoundRobinByAttributeRule.java:62-70), -1 = none.  [Q][N]                  */
  const int32_t* exec_id;       /* Offer.executor_ids, -1 pad.  [E][N]                                           */

  /* dynamicallyReservedPoolByResourceId (MesosResourcePool.java:119-131): resources carrying a resource-id label,
   * CSR by offer row.  All may be NULL when n_reserved == 0.                                                     */
  int64_t n_reserved;
  const int64_t* reserved_begin; /* [N+1]                                                                         */
  const int32_t* reserved_id;    /* resource-id dictionary id                                                     */
  const int32_t* reserved_kind;  /* scalar kind id, or OE_KIND_PORTS / OE_KIND_MOUNT                              */
  const double* reserved_value;  /* scalar amount | single port number | MOUNT size                              */
} oe_offer_table;

#define OE_KIND_PORTS 100
#define OE_KIND_MOUNT 101

/* ------------------------------------------------------------------------------------------------------------------
 * Task table — T rows = `stateStore.fetchTasks()` (OfferEvaluator.java:124), reduced to what placement rules read
 * back through S/offer/taskdata/TaskLabelReader.java:47-92.
 * ---------------------------------------------------------------------------------------------------------------- */
typedef struct {
  int64_t n_tasks;
  int32_t n_attr_slots;     /* TA */
  int32_t n_rr_cols;        /* Q (must equal the offer table's) */
  int32_t n_filter_words;   /* FW = ceil(n_filters / 32) */
  int32_t reserved0;
  const int32_t* name_id;       /* TaskInfo.name  [T]                                                            */
  const int32_t* pod_type_id;   /* label task_type, -1 = missing  [T]                                            */
  const int32_t* pod_index;     /* label index, OE_NO_INDEX = missing/unparsable  [T]                            */
  const int32_t* agent_id;      /* TaskInfo.slave_id  [T]                                                        */
  const int32_t* hostname_id;   /* label offer_hostname, -1 = missing (TaskException ⇒ no key)  [T]              */
  const int32_t* zone_id;       /* label offer_zone, -1 = missing  [T]                                           */
  const int32_t* region_id;     /* label offer_region, -1 = missing  [T]                                         */
  const int32_t* attr_id;       /* label offer_attributes split on ';', -1 pad.  [TA][T]                         */
  const int32_t* rr_key;        /* as oe_offer_table.rr_key, from the label strings.  [Q][T]                     */
  const uint32_t* filter_bits;  /* bit f set iff taskFilter_f.matches(TaskInfo.name).  [FW][T]                   */
} oe_task_table;

/* ------------------------------------------------------------------------------------------------------------------
 * Requirement classes: the stage pipeline (OfferEvaluator.java:250-310, 411-522, 538-728) of one pod shape,
 * flattened.  Instances of one pod type that are deployed fresh share a class.
 * ---------------------------------------------------------------------------------------------------------------- */
typedef enum {
  /* always-pass stages kept so that stage indices line up with the reference pipeline */
  OE_OP_NOP = 0,            /* LaunchEvaluationStage.java:42-90, TLSEvaluationStage.java:111, DestroyEvaluationStage  */
  /* ExecutorEvaluationStage.java:32-81: arg = executor id to reuse (OfferEvaluator.java:295-304) or -1; pass iff
   * arg < 0 or arg is listed in offer.exec_id.  A requirement that reuses an executor gets a class of its own.   */
  OE_OP_EXEC_ID = 1,
  /* PlacementRuleEvaluationStage.java:30-42: arg = first node, aux = node count of the postfix rule program     */
  OE_OP_RULE = 2,
  /* consumeReservableMerged (MesosResourcePool.java:330-372): d = amount - pool[slot][kind]; FAIL if the entry is
   * absent or d > 0; else pool -= amount.  Also VolumeEvaluationStage ROOT (VolumeEvaluationStage.java:155-181).  */
  OE_OP_SCALAR_NEW = 3,
  /* PortEvaluationStage.java:75-76,118-127 with a fixed host port: arg = port; FAIL unless free in slot's pool   */
  OE_OP_PORT_STATIC = 4,
  /* selectDynamicPort (PortEvaluationStage.java:242-298): lowest free port of the slot's pool that is not in the
   * class's `port_excl` list; arg = range-set id (-1 = unconstrained).  Chosen port reported in oe_decision.     */
  OE_OP_PORT_DYNAMIC = 5,
  /* consumeAtomic (MesosResourcePool.java:282-328): first unused MOUNT slot with amount - size <= 0 and
   * (profile >= 0 ? bit profile of `mask` : mask == 0).  Chosen slot reported in oe_decision.                    */
  OE_OP_VOL_MOUNT = 6,
  /* existing-pod path — evaluateSimpleResource with a resource id (OfferEvaluationUtils.java:60-227) on
   * consumeReserved (MesosResourcePool.java:245-280): arg = resource id.  FAIL if the offer lacks the id; equal ⇒
   * ok; amount > reserved ⇒ SCALAR_NEW on the difference; amount < reserved ⇒ ok (surplus unreserved by host).   */
  OE_OP_SCALAR_RESERVED = 7,
  OE_OP_PORT_RESERVED = 8,  /* same for a port kept from the prior launch: arg = resource id, aux = port          */
  OE_OP_MOUNT_RESERVED = 9, /* VolumeEvaluationStage.java:188-197: FAIL unless offer holds resource id `arg`      */
  /* UnreserveEvaluationStage.java:24-36 → MesosResourcePool.free (:374-430): the orphaned reservation `arg` goes back
   * into the pool and its id is forgotten; always passes.  Scalar kind: amount is added to pool[slot = previous
   * role][kind].  kind OE_KIND_PORTS: aux = the port, added to the slot's ports.  kind OE_KIND_MOUNT: amount = size,
   * aux = profile id + 1 (0 = none); the disk joins the end of the MOUNT list and is reported as mount slot
   * OE_MAX_MOUNT_SLOTS + k (k-th freed disk of the pipeline).                                                      */
  OE_OP_FREE = 10,
  OE_OP__COUNT
} oe_opcode;

/* Executor-level resource of an existing pod: when the executor id in `aux` is listed in the offer the stage passes
 * without touching the pool (ResourceEvaluationStage.java:65-84, VolumeEvaluationStage.java:126-154).              */
#define OE_OPF_SKIP_IF_RUNNING_EXECUTOR 0x01u

typedef struct {
  uint8_t opcode;   /* oe_opcode */
  uint8_t kind;     /* scalar kind id (OE_OP_FREE: also OE_KIND_PORTS / OE_KIND_MOUNT) */
  uint8_t slot;     /* role slot of spec.preReservedRole (OE_OP_FREE: of the resource's previous role) */
  uint8_t flags;    /* OE_OPF_* */
  int32_t arg;
  double amount;
  uint64_t mask;    /* VOL_MOUNT: allowed-profile mask */
  uint64_t aux;
} oe_stage_op;      /* 32 bytes */

typedef enum {
  OE_R_TRUE = 0,          /* P/PassthroughRule.java:25                                                            */
  OE_R_FALSE = 1,         /* P/InvalidPlacementRule.java:34                                                       */
  OE_R_AND = 2,           /* a = child count; zero children ⇒ FAIL (P/AndRule.java:40-44)                         */
  OE_R_OR = 3,            /* a = child count; zero children ⇒ FAIL (P/OrRule.java:36-58)                          */
  OE_R_NOT = 4,           /* P/NotRule.java:28-37                                                                 */
  OE_R_AGENT_EQ = 5,      /* a = agent id (P/AgentRule.java:98-115)                                               */
  OE_R_ATTR_ANY = 6,      /* a = matcher_bits word offset over the attribute dictionary (P/AttributeRule.java)    */
  OE_R_HOST_IN = 7,       /* a = word offset over the hostname dictionary (P/HostnameRule.java:26-52)             */
  OE_R_ZONE_IN = 8,       /* a = word offset over zones; no fault domain ⇒ FAIL (P/ZoneRule.java:39-40)           */
  OE_R_REGION_IN = 9,     /* a = word offset over regions; no fault domain ⇒ FAIL (P/RegionRule.java:25-51)       */
  OE_R_LOCAL_REGION = 10, /* a = local region id, b = 1 iff the local domain is known (P/IsLocalRegionRule.java)  */
  OE_R_MAXPER_HOST = 11,  /* a = counter set, b = max (P/MaxPerHostnameRule.java:112-124)                         */
  OE_R_MAXPER_ZONE = 12,  /* a, b as above; no fault domain ⇒ FAIL (P/MaxPerZoneRule.java:58-59)                  */
  OE_R_MAXPER_REGION = 13,/* a, b as above; no fault domain ⇒ PASS (no keys, P/MaxPerRegionRule.java:33-45)       */
  OE_R_MAXPER_ATTR = 14,  /* a = counter set, b = max, c = word offset of the attribute matcher over attrs
                             (P/MaxPerAttributeRule.java:137-153)                                                 */
  OE_R_RR = 15,           /* a = counter set (its keyspace selects hostname/zone/region/attribute column),
                             b = distinct key count or -1 (P/AbstractRoundRobinRule.java:53-151)                  */
  OE_R_TYPE_AVOID = 16,   /* a = type set (P/TaskTypeRule.java:98-175)                                            */
  OE_R_TYPE_COLOCATE = 17,/* a = type set (P/TaskTypeRule.java:177-200)                                           */
  OE_R__COUNT
} oe_rule_op;

typedef struct {
  int32_t op; /* oe_rule_op */
  int32_t a, b, c;
} oe_rule_node; /* 16 bytes */

typedef enum {
  OE_KS_HOSTNAME = 0, /* key = hostname_id (MaxPerHostname / RoundRobinByHostname)                                */
  OE_KS_ZONE = 1,
  OE_KS_REGION = 2,
  OE_KS_ATTR = 3,     /* multi-key: every attribute string of the row (MaxPerAttribute)                           */
  OE_KS_RR_ATTR = 4   /* key = rr_key[col] (RoundRobinByAttribute)                                                */
} oe_keyspace;

/* One (key space, task filter) pair: the per-key tallies that MaxPerRule.java:52-87 and
 * AbstractRoundRobinRule.java:70-89 rebuild from all tasks for every offer.  Reduced once per cycle on the device. */
typedef struct {
  int32_t keyspace;  /* oe_keyspace */
  int32_t rr_col;    /* column for OE_KS_RR_ATTR, else 0 */
  int32_t filter_id; /* bit index into oe_task_table.filter_bits */
  int32_t reserved0;
} oe_counter_set;

typedef struct {
  int32_t n_classes;
  int32_t n_ops;
  int32_t n_rule_nodes;
  int32_t n_counter_sets;
  int32_t n_type_sets;
  int32_t n_range_sets;
  int64_t n_matcher_words;
  const int32_t* op_begin;          /* [C+1] into ops                                                             */
  const oe_stage_op* ops;           /* [n_ops]                                                                    */
  const oe_rule_node* rule_nodes;   /* [n_rule_nodes] postfix programs, addressed by OE_OP_RULE                   */
  /* OfferEvaluationUtils.getRole(podSpec) (OfferEvaluationUtils.java:229-235): role id or -1 when absent  [C]   */
  const int32_t* pod_role;
  /* static host ports declared anywhere in the pod (PortEvaluationStage.java:252-261), sorted ascending          */
  const int32_t* port_excl_begin;   /* [C+1]                                                                      */
  const uint16_t* port_excl;
  /* PortSpec.getRanges() of dynamic ports (S/specification/RangeSpec.java:12-54), packed like port_range         */
  const int32_t* range_set_begin;   /* [n_range_sets+1]                                                           */
  const uint32_t* range_set;
  const oe_counter_set* counter_sets; /* [n_counter_sets]                                                         */
  const int32_t* type_set_pod_type;   /* [n_type_sets] pod type id each TaskTypeRule looks for                     */
  const uint32_t* matcher_bits;       /* [n_matcher_words] bit v of a matcher = matcher.matches(dictionary[v])    */
  oe_dict_sizes dict;
} oe_program_table;

/* ------------------------------------------------------------------------------------------------------------------
 * Requirement table — the candidate steps of one cycle in PlanScheduler order
 * (S/scheduler/plan/PlanScheduler.java:56-67); steps that are not pending/prepared (:73-78) are left out by the host.
 * ---------------------------------------------------------------------------------------------------------------- */
typedef struct {
  int64_t n_reqs;
  const int32_t* class_id;     /* [M] */
  const int32_t* pod_type_id;  /* PodInstance.getPod().getType()  [M] */
  const int32_t* pod_index;    /* PodInstance.getIndex()  [M] */
  /* deployed tasks equivalent to the requirement (S/offer/TaskUtils.java:136-145) as rows of the task table: the
   * only per-instance input of the placement rules.  CSR  [M+1]; NULL ⇒ the engine derives it from the task table */
  const int32_t* self_begin;
  const int32_t* self_rows;
  /* multi-cycle replay (oe_apply_launches): the tasks a launch of this pod stores
   * (LaunchEvaluationStage.java:49-68 → S/state/PersistentLaunchRecorder.java:74-137), CSR [M+1]; may be NULL.   */
  const int32_t* launch_begin;
  const int32_t* launch_name_id;      /* task name id per stored task                                            */
  const uint32_t* launch_filter_bits; /* [n_launch_tasks][FW] row-major: filter bits of each stored task's name  */
} oe_req_table;

/* ------------------------------------------------------------------------------------------------------------------
 * Results
 * ---------------------------------------------------------------------------------------------------------------- */
typedef struct {
  int32_t offer_row;                    /* GLOBAL list index of the accepted offer, -1 = no offer fits            */
  int32_t n_new_reservations;           /* RESERVE recommendations that need a fresh UUID from the host
                                           (S/offer/ResourceBuilder.java:318,343), in stage order                 */
  uint16_t dyn_port[OE_MAX_DYN_PORTS];  /* port chosen by the i-th PORT_DYNAMIC op (0 = unused)                   */
  int8_t mount_slot[OE_MAX_MOUNT_OPS];  /* MOUNT slot chosen by the i-th VOL_MOUNT op (-1 = unused)               */
  uint32_t reserved0;
} oe_decision; /* 32 bytes */

typedef struct {
  float ms_h2d;          /* host→device copies issued by the last evaluate call (0 when tables were resident)    */
  float ms_counters;     /* K0: task table → per-key tallies                                                      */
  float ms_feasibility;  /* K1: offers × classes stage programs + rule programs.  While an OE_CFG_ASYNC_LOADS offer
                            table is still arriving this span includes the waits for its chunks (the bus), and
                            ms_commit is only what remains of the commit after the last chunk's K1: the pass over
                            the landed chunks ran beside the bus                                                  */
  float ms_expand;       /* K1x: class bits → offers × requirements bit matrix (in oe_evaluate_cycle it runs BESIDE
                            the commit: ms_expand and ms_commit + ms_d2h then overlap)                            */
  float ms_commit;       /* K2: ordered greedy commit + winner materialisation                                    */
  float ms_d2h;          /* device→host copy of the results                                                       */
  float ms_total;
  int32_t n_placed;      /* requirements that got an offer                                                        */
  int64_t n_pairs;       /* offer × requirement pairs decided                                                     */
  int64_t kernel_launches; /* kernels launched by the call                                                        */
} oe_stats;

/* ------------------------------------------------------------------------------------------------------------------
 * Loading.  Order: programs, tasks, offers, requirements (each may be re-loaded between cycles).  Re-loading programs
 * invalidates tasks and requirements.  The RESIDENT OFFER TABLE stays valid across a program re-load when every dictionary
 * it indexes (agents, hostnames, zones, regions, attributes, round-robin values, executor ids, resource ids) is at least as
 * large as before — ids are never re-numbered by a growing dictionary — and the new stage ops name only role slots, scalar
 * kinds and round-robin columns the table has; otherwise the next call that needs offers fails with OE_ERR_STATE until
 * oe_load_offers ran again.  That is what lets an adapter keep the table on the device from cycle to cycle and send only the
 * offers that changed (oe_update_offers) although the plan, and with it the classes, changes every cycle.
 * ---------------------------------------------------------------------------------------------------------------- */
int32_t oe_load_programs(oe_engine* h, const oe_program_table* t);
int32_t oe_load_tasks(oe_engine* h, const oe_task_table* t);
int32_t oe_load_offers(oe_engine* h, const oe_offer_table* t);
/* Delta upload: replaces rows of the RESIDENT offer table in place.  Between two cycles only a few agents' offers change
 * (the reference's queue takes what arrived and drops what was rescinded, S/framework/OfferQueue.java:46-85, and the
 * evaluator sees the new list, S/framework/OfferProcessor.java:240-262); the host re-encodes those and sends them here
 * instead of the whole table.  Row rows[k] (local row index of this engine, 0 <= rows[k] < n_offers, all distinct) receives
 * row k of `values`, a table of n_rows rows with the same slot counts as the resident one (n_offers = n_rows; row_offset,
 * n_offers_global and shard_rows are ignored).  A withdrawn offer is a row whose scalar_present is 0 and whose port ranges
 * are all OE_PORT_EMPTY: nothing fits it.  Tables with a reserved side table (resident or in `values`) cannot be patched —
 * OE_ERR_STATE, reload with oe_load_offers.  Dictionary ids are validated like those of a load.  `values` is always at full
 * width, also when the resident table was loaded with oe_load_offers_packed: the engine looks the replaced rows' `ports` range
 * lists up in the resident list dictionary and appends the lists nobody had (room for max(4096, a quarter) more than were
 * loaded; beyond that the feasibility pass goes back to per-offer port tests until the table is loaded whole again).        */
int32_t oe_update_offers(oe_engine* h, const int64_t* rows, int64_t n_rows, const oe_offer_table* values);
int32_t oe_load_requirements(oe_engine* h, const oe_req_table* t);

/* Compact wire format of the offer table.  An offer row of oe_offer_table is ~100 bytes, and at 10^6 offers the PCIe copy of
 * the table IS the placement cycle's latency; most of those bytes carry no information: the allocation role is the same for
 * every row, zone / region / attribute ids fit one byte, the scalar amounts Mesos reports are exact in fp32, and nearly all
 * agents offer one of a handful of `ports` range lists.  oe_load_offers_packed takes the same oe_offer_table, but its column
 * pointers are read as this struct says; the engine widens the columns on the device (exactly: no value changes) into the
 * layout oe_load_offers builds, so every evaluation call behaves as after a plain load.  The encoder decides per column and
 * per table (the host side of S/offer/MesosResourcePool.java:76-147: it sees every value while it sums the pools).
 *   *_bytes           element size on the wire of that id column: 4 (int32 as in oe_offer_table), 2 or 1 (unsigned; the type's
 *                     all-ones value stands for -1, so the dictionary must be smaller than that)
 *   scalar_is_f32     `scalar` points at float [S][K][N]; the encoder verified value == (double)(float)value for every entry
 *   hostname_is_agent hostname_id is not read: every row's hostname id equals its agent id (one hostname per agent, ids
 *                     assigned together)
 *   alloc_role_is_const  alloc_role is not read: every row holds alloc_role_const (-1 when no offer carries an allocation role)
 *   port_tuple_bytes  0: port_range is sent per offer as in oe_load_offers.  2 or 4: port_range is NOT read; port_tuples holds
 *                     the n_port_tuples DISTINCT range lists of the table ([n_port_tuples][R], each sorted, disjoint,
 *                     OE_PORT_EMPTY padded) and port_tuple_id [S][N] (unsigned, that many bytes) names each (slot, offer)'s list.
 *                     The feasibility kernel then evaluates a class's port stages once per distinct list instead of once per offer.
 * Everything else (row counts, slot counts, mount_size, agent_id, the reserved side table) is as in oe_offer_table.           */
typedef struct {
  uint8_t scalar_is_f32;
  uint8_t present_bytes;
  uint8_t zone_bytes, region_bytes, attr_bytes, rr_bytes, mount_profile_bytes, exec_bytes;
  uint8_t hostname_is_agent;
  uint8_t alloc_role_is_const;
  uint8_t port_tuple_bytes;
  uint8_t reserved0[5];
  int32_t alloc_role_const;
  int32_t n_port_tuples;
  const uint32_t* port_tuples;
  const void* port_tuple_id;
} oe_offer_packing;
int32_t oe_load_offers_packed(oe_engine* h, const oe_offer_table* t, const oe_offer_packing* pk);

/* ------------------------------------------------------------------------------------------------------------------
 * Evaluation
 * ---------------------------------------------------------------------------------------------------------------- */
/* Row stride, in 32-bit words, of the bit matrices below: ceil(n_offers / 32) rounded up to a multiple of 128
 * (512-byte row alignment for 128-bit stores); with world > 1 it is derived from shard_rows so every shard agrees.     */
int64_t oe_matrix_stride_words(const oe_engine* h);

/* Full feasibility matrix of the loaded tables: bit (i & 31) of word [j * stride + (i >> 5)] is 1 iff ALL stages of
 * requirement j pass on offer row i — the boolean that OfferEvaluator.java:165-184 computes for one (requirement,
 * offer).  Runs K0 + K1 + K1x on the device; when `bits` is non-NULL the matrix (n_reqs * stride words) is copied to
 * it.  Placement tallies are those of the loaded task table (frozen, as within one reference cycle).              */
int32_t oe_feasibility(oe_engine* h, uint32_t* bits, oe_stats* stats);

/* Class-factored matrix: one row per requirement class (fresh instances) — what the commit kernel scans.          */
int32_t oe_class_feasibility(oe_engine* h, uint32_t* bits, oe_stats* stats);

/* Which row of the class-factored matrix (oe_class_feasibility, OE_BUF_CLASS_BITS_*) each requirement reads, as of the last
 * feasibility pass on this handle: its class's row, or — when the requirement's pod is already deployed (a stored task of the
 * same pod type and index, S/offer/TaskUtils.java:136-145) and its class's rule reads tallies — a row behind the classes that
 * is evaluated without the pod's own tasks (P/MaxPerRule.java:60-62).  Such requirements share a row when their rule verdicts
 * are identical.  rows: [n_reqs] (may be NULL); *n_rows: rows of the matrix.  OE_ERR_STATE before the first pass.          */
int32_t oe_requirement_rows(oe_engine* h, int32_t* rows, int32_t* n_rows);

/* One PlanScheduler.resourceOffers cycle (S/scheduler/plan/PlanScheduler.java:50-70): requirements in order, each
 * takes the FIRST offer in list order whose stages all pass (OfferEvaluator.java:137-245) among the offers not yet
 * taken by an earlier requirement of this cycle (:61-66).  out[M].  When world > 1 the caller must run
 * oe_class_feasibility() on every shard, gather the shards' rows (see oe_device_buffer) and then call this on every
 * rank: all ranks compute the same decisions.  With peers attached (oe_peer_attach) the gather is part of the pass
 * and this call alone is the sharded cycle.                                                                        */
int32_t oe_evaluate_cycle(oe_engine* h, oe_decision* out, oe_stats* stats);

/* Why did an offer not fit?  For requirements [req_begin, req_end) and every loaded offer: bit min(s, 63) of
 * failed[(j - req_begin) * n_offers + i] is set iff stage s of requirement j's pipeline fails on offer i.  All stages are
 * run, as the reference does (S/offer/evaluate/OfferEvaluator.java:165-171: no short-circuit; a failed stage leaves the pool
 * untouched and later stages still run), so the masks are what OfferEvaluator.java:173-214 records into the offer-outcome
 * trackers (S/debug/OfferOutcomeTrackerV2.java:52-120, a ring of the 100 most recent offers).  Mask 0 = the pair is feasible. */
int32_t oe_stage_failures(oe_engine* h, int64_t req_begin, int64_t req_end, uint64_t* failed);

/* Reservation garbage scan — DefaultScheduler.getUnexpectedResources (S/scheduler/DefaultScheduler.java:483-538): every
 * reserved resource (an entry of the offer table's reserved side table, i.e. one that carries a resource-id label) whose id
 * is NOT in the keep set is unexpected and will be unreserved / destroyed by the host.  keep_bits: bit r set iff resource id
 * r belongs to a stored task that is neither permanently failed nor being decommissioned (:487-499), ceil(n_res_ids / 32)
 * words.  With only_unused_offers != 0 the entries of offers accepted by the last oe_evaluate_cycle are skipped (the
 * reference scans `unusedOffers`, S/framework/OfferProcessor.java:451-455).  unexpected[q] (one byte per side-table entry,
 * in table order) receives 0 / 1; *n_unexpected the number of ones.                                                   */
int32_t oe_unexpected_resources(oe_engine* h, const uint32_t* keep_bits, int32_t only_unused_offers, uint8_t* unexpected,
                                int64_t* n_unexpected);

/* Multi-cycle replay: store the tasks of every pod placed by the last oe_evaluate_cycle into the device task table
 * (rows with the same task name are overwritten, as StateStore.storeTasks does), so that the next cycle's tallies
 * see them (LaunchEvaluationStage.java:58-68 → PersistentLaunchRecorder.java:74-137).  On a sharded engine: see
 * oe_gather_decisions; every rank calls it after its oe_evaluate_cycle and ends up with the same task table.         */
int32_t oe_apply_launches(oe_engine* h);

/* Current device task table (after oe_apply_launches), slot-major columns sized by *n_tasks; pass name_id == NULL to
 * query the row count only.                                                                                        */
int32_t oe_export_tasks(oe_engine* h, int64_t* n_tasks, int32_t* name_id, int32_t* pod_type_id, int32_t* pod_index,
                        int32_t* agent_id, int32_t* hostname_id, int32_t* zone_id, int32_t* region_id, int32_t* attr_id,
                        int32_t* rr_key, uint32_t* filter_bits);

/* Device-resident buffers, for callers that keep data on the GPU (benchmarks, multi-GPU gather).                   */
typedef enum {
  OE_BUF_CLASS_BITS_LOCAL = 0,  /* [n_rows_factored][local stride] this shard's class-factored rows               */
  OE_BUF_CLASS_BITS_GLOBAL = 1, /* [world][n_rows_factored][local stride] gathered rows, shard-major              */
  OE_BUF_MATRIX = 2,            /* [n_reqs][stride] full matrix of this shard                                      */
  OE_BUF_DECISIONS = 3          /* [n_reqs] oe_decision                                                            */
} oe_buffer_id;
int32_t oe_device_buffer(oe_engine* h, int32_t which, void** dev_ptr, int64_t* n_bytes);
/* The stream the engine launches on (cudaStream_t), so that callers can order their own work / events after it.   */
void* oe_stream(oe_engine* h);

/* ------------------------------------------------------------------------------------------------------------------
 * Peer exchange for sharded engines (world > 1, one engine per GPU of an NVLink box).  The reference evaluates every
 * requirement against the ONE list of all offers (S/scheduler/plan/PlanScheduler.java:56-66: `availableOffers`), so a
 * sharded cycle has one exchange step: every rank needs every shard's class-factored rows before it replays the commit.
 * With peers attached that step is part of the feasibility pass: the kernel stores each 16 bytes of rows it produces
 * straight into the exchange region of EVERY shard (its own included) through NVLink peer memory, raises a flag per
 * shard behind them, and oe_evaluate_cycle waits on the device for all shards' flags before the commit reads its own
 * region.  No collective library call and no host round trip sit between the pass and the commit, and
 * oe_evaluate_cycle alone is then the whole sharded cycle (it runs the pass itself when the class rows are stale).
 *
 *   oe_peer_export  after oe_load_offers (the row stride comes from shard_rows): allocates this shard's region for up
 *                   to max_rows factored rows (classes + variants; a pass with more fails with OE_ERR_LIMIT — export
 *                   and attach again with a larger bound), resets the pass counter, and returns the region's CUDA IPC
 *                   handle (OE_PEER_HANDLE_BYTES, may be NULL) and its device address (may be NULL).
 *   oe_peer_attach  handles: world x OE_PEER_HANDLE_BYTES in rank order, from every rank's oe_peer_export (NULL when
 *                   all shards are given as pointers); same_process_ptrs: NULL, or world entries, non-NULL where the
 *                   shard's engine lives in THIS process (an IPC handle cannot be opened by the process that made it).
 * Both are collective in the sense that every rank calls them, in this order, and nobody evaluates before all have
 * attached.  All ranks must then issue the same sequence of evaluation calls; a rank may publish at most one pass ahead
 * of a peer that is still committing (two buffers alternate), which oe_evaluate_cycle guarantees by waiting — so call
 * it (not only oe_class_feasibility) once per pass.  A shard that does not publish within 10 s (OE_PEER_TIMEOUT_MS in
 * the environment at export time) makes oe_evaluate_cycle fail with OE_ERR_STATE instead of waiting forever.  At most
 * 16 shards.  Without peers attached the gather stays the caller's job (oe_device_buffer).                           */
#define OE_PEER_HANDLE_BYTES 64
int32_t oe_peer_export(oe_engine* h, int32_t max_rows, void* handle, void** dev_ptr);
int32_t oe_peer_attach(oe_engine* h, const void* handles, void* const* same_process_ptrs);
/* Back to the caller-side gather: unmaps the other shards' regions (when one rank could not attach, all ranks detach). */
int32_t oe_peer_detach(oe_engine* h);
/* The complete decision table of a sharded cycle, in one place.  Every rank computes every offer_row, but the dynamic ports,
 * MOUNT slots and the number of fresh reservation ids of a winner (S/offer/evaluate/PortEvaluationStage.java:242-298,
 * S/offer/MesosResourcePool.java:282-328) can only be filled in by the shard that holds the winning offer.  With peers attached
 * and a decision area in the exchange regions (oe_peer_export AFTER the requirement table was loaded: the area is sized for
 * it, plus a quarter), oe_evaluate_cycle stores each shard's winners into every shard's region; this call waits on the device
 * until all shards' records of the pass have landed and writes the merged table to out[M] (may be NULL: only merge on the
 * device).  Call it on the ranks that want the table — oe_evaluate_cycle itself accepts out == NULL on a peer-attached engine,
 * so that only one rank pays for the copy.  oe_apply_launches on a sharded engine does this step itself and reads the winning
 * offers' placement columns from the same records (S/state/PersistentLaunchRecorder.java:74-137 has ONE task store).       */
int32_t oe_gather_decisions(oe_engine* h, oe_decision* out);

/* ------------------------------------------------------------------------------------------------------------------
 * Allocation-free first-fit.  This entry point has no counterpart in the reference (the JVM allocates freely); it
 * exists for hosts that cannot let this library call cudaMalloc/cudaFree on their behalf: a multi-service scheduler
 * running several engines on one GPU (an allocation or free synchronises the whole device and would stall the other
 * services' cycles), or a process whose device memory is owned up front by a pool or capped by an operator budget.
 *
 * Computes the `offer_row` column of oe_evaluate_cycle — requirements in order, each takes the lowest-index offer
 * whose bit is set in its row and that no earlier requirement took (S/scheduler/plan/PlanScheduler.java:56-67) —
 * from bit rows, a requirement→row map and into a result column that all live in DEVICE memory owned by the caller:
 *   d_row_bits  [n_rows][stride_words]  bit (i & 31) of word [r * stride_words + (i >> 5)] = offer i fits row r
 *                                       (the layout of OE_BUF_CLASS_BITS_LOCAL)
 *   d_req_row   [n_reqs]                bit row of each requirement, in plan order (out-of-range row ⇒ no offer)
 *   d_offer_row [n_reqs]                result: offer index, or -1 when nothing is left for that requirement
 * The kernel is enqueued on `stream` (a cudaStream_t, NULL = default stream) of `device` and the call returns without
 * waiting for it.  No memory is allocated or freed, on the device or the host.  The working memory (one "taken" bit
 * per offer, one cursor per row) is a fixed array compiled into the library, held by every GPU that loads it until
 * the process exits (~544 KiB); hence the limits below, beyond which the call returns OE_ERR_LIMIT with an
 * explanation in err_text.  The first call on a given device also looks up that array's device address; make that
 * first call outside any stream capture.  All calls on one device share the array: the kernel re-zeroes what it uses,
 * but two calls whose kernels overlap on the GPU (different streams, or different host threads) corrupt each other —
 * serialise them.  err_text (err_cap bytes, may be NULL) receives the text of a failure.                              */
#define OE_STATIC_MAX_OFFERS (4 * 1024 * 1024)
#define OE_STATIC_MAX_ROWS 8192
int32_t oe_first_fit_static(int32_t device, void* stream, const uint32_t* d_row_bits, int64_t stride_words, int32_t n_rows,
                            int64_t n_offers, const int32_t* d_req_row, int64_t n_reqs, int32_t* d_offer_row,
                            char* err_text, int32_t err_cap);

#ifdef __cplusplus
}
#endif
#endif /* OE_ABI_H_ */
