/*
 * ref_pmm_glue.c -- TEST INFRASTRUCTURE ONLY (oracle/_ref build -> libref_pmm.so).
 *
 * Compiles the reference's eviction-order code VERBATIM -- line-range extracts of kernel-open/nvidia-uvm/uvm_pmm_gpu.c:
 *   :1395-1414  chunk_is_evictable                 :1416-1428  chunk_start_eviction
 *   :1430-1448  root_chunk_update_eviction_list    :1450-1458  uvm_pmm_gpu_mark_root_chunk_used / _unused
 *   :1460-1500  pick_root_chunk_to_evict
 * against a minimal PMM: root chunks with a list node, the two eviction lists (root_chunks.va_block_used /
 * va_block_unused) and two empty free lists (the allocator only evicts when it has nothing free).  What the
 * reference does around these functions is replayed by the harness below as a sequence of events per va_block
 * (= per 2 MB root chunk, uvm_va_block.c:3547-3566):
 *   ALLOC   uvm_pmm_gpu_unpin_allocated -> chunk_update_lists_locked (:627-651): tail of va_block_used
 *   USED    block_mark_memory_used (uvm_va_block.c:3547-3566)           -> uvm_pmm_gpu_mark_root_chunk_used
 *   UNUSED  block_clear_resident_processor (uvm_va_block.c:3579-3603)   -> uvm_pmm_gpu_mark_root_chunk_unused
 *   PICK    pick_root_chunk_to_evict; the victim is evicted and its chunk freed (evict_root_chunk :1290-1380)
 *
 * What this pins: N1's victim order -- unused chunks before used ones, each list in the order of its last move.
 */
#include "shim/kshim.h"
#include "ref_api.h"

/* list operations the shim does not have (include/linux/list.h semantics) */
static inline void list_del_init(struct list_head *e) { e->prev->next = e->next; e->next->prev = e->prev; INIT_LIST_HEAD(e); }
static inline void list_move_tail(struct list_head *e, struct list_head *head)
{ e->prev->next = e->next; e->next->prev = e->prev; list_add_tail(e, head); }
#define list_first_entry_or_null(head, type, member) \
    (list_empty(head) ? (type *)NULL : (type *)((char *)(head)->next - offsetof(type, member)))

/* ---- the PMM fields the extracts touch (uvm_pmm_gpu.h:232-330, 340-420) ---- */
typedef enum {
    UVM_PMM_GPU_CHUNK_STATE_PMA_OWNED, UVM_PMM_GPU_CHUNK_STATE_FREE, UVM_PMM_GPU_CHUNK_STATE_IS_SPLIT,
    UVM_PMM_GPU_CHUNK_STATE_TEMP_PINNED, UVM_PMM_GPU_CHUNK_STATE_ALLOCATED
} uvm_pmm_gpu_chunk_state_t;
typedef enum { UVM_PMM_GPU_MEMORY_TYPE_USER, UVM_PMM_GPU_MEMORY_TYPE_KERNEL } uvm_pmm_gpu_memory_type_t;
typedef enum { UVM_PMM_LIST_NO_ZERO, UVM_PMM_LIST_ZERO } uvm_pmm_list_zero_t;
#define UVM_CHUNK_SIZE_MAX (2u << 20)
typedef struct uvm_gpu_chunk_struct {
    struct list_head list;
    uvm_pmm_gpu_chunk_state_t state;
    bool is_zero, in_eviction;
    uint32_t slot;
} uvm_gpu_chunk_t;
typedef struct { uvm_gpu_chunk_t chunk; } uvm_gpu_root_chunk_t;
typedef struct {
    int list_lock;
    struct { struct list_head va_block_used, va_block_unused; } root_chunks;
    struct list_head free_list[2];
} uvm_pmm_gpu_t;
#define uvm_spin_lock(l) ((void)(l))
#define uvm_spin_unlock(l) ((void)(l))
#define uvm_assert_spinlock_locked(l) ((void)(l))
static uvm_gpu_root_chunk_t *root_chunk_from_chunk(uvm_pmm_gpu_t *pmm, uvm_gpu_chunk_t *chunk) { (void)pmm; return (uvm_gpu_root_chunk_t *)chunk; }
static bool chunk_is_root_chunk_pinned(uvm_pmm_gpu_t *pmm, uvm_gpu_chunk_t *chunk) { (void)pmm; (void)chunk; return false; }
static bool chunk_is_in_eviction(uvm_pmm_gpu_t *pmm, uvm_gpu_chunk_t *chunk) { (void)pmm; return chunk->in_eviction; }
static void uvm_gpu_chunk_set_in_eviction(uvm_gpu_chunk_t *chunk, bool v) { chunk->in_eviction = v; }
static NvU32 uvm_gpu_chunk_get_size(uvm_gpu_chunk_t *chunk) { (void)chunk; return UVM_CHUNK_SIZE_MAX; }
static bool uvm_gpu_chunk_is_user(const uvm_gpu_chunk_t *chunk) { (void)chunk; return true; }
static struct list_head *find_free_list(uvm_pmm_gpu_t *pmm, uvm_pmm_gpu_memory_type_t type, NvU32 size, uvm_pmm_list_zero_t z)
{ (void)type; (void)size; return &pmm->free_list[z]; }
static uvm_gpu_chunk_t *list_first_chunk(struct list_head *list) { return list_first_entry_or_null(list, uvm_gpu_chunk_t, list); }

#include "x_pmm.h"              /* the uvm_pmm_gpu.c extracts listed at the top, verbatim */

/* ======================= flat harness API ======================= */
enum { EV_ALLOC = 0, EV_USED = 1, EV_UNUSED = 2, EV_PICK = 3 };

/* Replays n events; events[i] = {kind, slot}.  PICK events write the victim's slot (0xFFFFFFFF: nothing to evict)
 * to victims[] in order and free the victim's chunk.  Returns the number of victims. */
uint32_t ref_pmm_replay(uint32_t n_slots, uint32_t n_events, const uint32_t *events, uint32_t *victims)
{
    uvm_pmm_gpu_t pmm;
    uvm_gpu_root_chunk_t *rc = (uvm_gpu_root_chunk_t *)calloc(n_slots, sizeof(*rc));
    uint32_t i, nv = 0;
    INIT_LIST_HEAD(&pmm.root_chunks.va_block_used);
    INIT_LIST_HEAD(&pmm.root_chunks.va_block_unused);
    INIT_LIST_HEAD(&pmm.free_list[0]);
    INIT_LIST_HEAD(&pmm.free_list[1]);
    for (i = 0; i < n_slots; i++) {
        INIT_LIST_HEAD(&rc[i].chunk.list);
        rc[i].chunk.state = UVM_PMM_GPU_CHUNK_STATE_FREE;
        rc[i].chunk.slot = i;
    }
    for (i = 0; i < n_events; i++) {
        const uint32_t kind = events[2 * i], slot = events[2 * i + 1];
        uvm_gpu_chunk_t *c = slot < n_slots ? &rc[slot].chunk : NULL;
        if (kind == EV_ALLOC) {
            /* chunk_update_lists_locked :639-643 for an ALLOCATED root chunk */
            c->state = UVM_PMM_GPU_CHUNK_STATE_ALLOCATED;
            list_move_tail(&c->list, &pmm.root_chunks.va_block_used);
        }
        else if (kind == EV_USED) {
            if (c->state == UVM_PMM_GPU_CHUNK_STATE_ALLOCATED)
                uvm_pmm_gpu_mark_root_chunk_used(&pmm, c);
        }
        else if (kind == EV_UNUSED) {
            if (c->state == UVM_PMM_GPU_CHUNK_STATE_ALLOCATED)
                uvm_pmm_gpu_mark_root_chunk_unused(&pmm, c);
        }
        else {
            uvm_gpu_root_chunk_t *v = pick_root_chunk_to_evict(&pmm);
            victims[nv++] = v ? v->chunk.slot : 0xFFFFFFFFu;
            if (v) {                                     /* evicted and freed (evict_root_chunk :1290-1380) */
                uvm_gpu_chunk_set_in_eviction(&v->chunk, false);
                v->chunk.state = UVM_PMM_GPU_CHUNK_STATE_FREE;
                INIT_LIST_HEAD(&v->chunk.list);
            }
        }
    }
    free(rc);
    return nv;
}
