#!/usr/bin/env bash
# build_ref.sh -- TEST INFRASTRUCTURE ONLY.
# Compiles the reference's own C for the hot path, from the sources where they lie under
# /root/reference, against the thin shims in oracle/shim/.  Outputs ONLY shared objects into
# oracle/_ref/ (git-ignored, but shipped to the GPU box).  Line-range extracts of reference
# headers are written to a temp dir and deleted; reference sources never enter the repo.
#
# Every extract is pinned by the sha256 of its source file: if the reference
# changes, the build fails instead of silently extracting the wrong lines.
set -euo pipefail

REF="${UVMTIER_REFERENCE:-/root/reference}"
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/_ref"
UVM="$REF/kernel-open/nvidia-uvm"
INC="$REF/kernel-open/common/inc"

if [ ! -d "$UVM" ]; then
    echo "build_ref: $REF not present; keeping prebuilt oracle/_ref/*.so" >&2
    exit 0
fi

pin() { # file sha256-prefix
    local got
    got="$(sha256sum "$UVM/$1" | cut -c1-16)"
    if [ "$got" != "$2" ]; then
        echo "build_ref: $1 sha256 prefix $got != pinned $2" >&2
        exit 1
    fi
}
pin uvm_perf_prefetch.c fe3d4e338321cecb
pin uvm_perf_prefetch.h cfc4f49f47d627ef
pin uvm_perf_utils.h 975610edc6739fea
pin uvm_va_block.c 085ac82080044437
pin uvm_va_block.h f1d0ce29432b4487
pin uvm_va_block_types.h 5c35bec1c234c184
pin uvm_gpu_replayable_faults.c c7991f9a9e029ae1
pin uvm_hal_types.h 5b01cb3403d43919
pin uvm_volta_fault_buffer.c cb8333be2239b583
pin clc369.h 755c9258db3cbce8

TMP="$(mktemp -d /tmp/uvmtier_ref.XXXXXX)"
trap 'rm -rf "$TMP"' EXIT
mkdir -p "$OUT"

CFLAGS="-std=gnu11 -O2 -fPIC -shared -Wall -Wno-unused-function -Wno-unused-variable -Wno-unused-but-set-variable"

# ---------------- prefetch leg: uvm_perf_prefetch.c, whole file, unmodified ----------------
sed -n '36,80p' "$UVM/uvm_va_block_types.h" > "$TMP/x_va_block_types.h"
sed -n '1604,1704p;1706,1882p;1915,1995p;2008,2039p' "$UVM/uvm_va_block.h" > "$TMP/x_va_block_helpers.h"
sed -n '2177,2189p;2202,2220p' "$UVM/uvm_va_block.c" > "$TMP/x_big_page_region.h"
gcc $CFLAGS -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_prefetch_glue.c" -o "$OUT/libref_prefetch.so"
echo "build_ref: built $OUT/libref_prefetch.so"

# ---------------- fault leg ----------------
if [ -f "$HERE/ref_faults_glue.c" ]; then
    sed -n '33,51p;76,98p;201,305p;309,317p;319,325p;342,453p' "$UVM/uvm_hal_types.h" > "$TMP/x_hal_types.h"
    sed -n '58,78p' "$UVM/uvm_push_macros.h" > "$TMP/x_push_macros.h"
    sed -n '105,152p;199,243p;312,391p' "$UVM/uvm_volta_fault_buffer.c" > "$TMP/x_volta_parse.h"
    sed -n '695,740p;742,986p;988,1020p;1118,1212p' "$UVM/uvm_gpu_replayable_faults.c" > "$TMP/x_faults.h"
    sed -n '57,83p' "$UVM/uvm_blackwell_fault_buffer.h" > "$TMP/x_blackwell_utlb_enum.h"
    sed -n '128,201p' "$UVM/uvm_blackwell_mmu.c" > "$TMP/x_blackwell_client_to_utlb.h"
    gcc $CFLAGS -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_faults_glue.c" "$HERE/ref_faults_blackwell.c" \
        "$HERE/ref_faults_hopper.c" -o "$OUT/libref_faults.so"
    echo "build_ref: built $OUT/libref_faults.so"
fi

# ---------------- service leg: service_fault_batch_block_locked and its helpers, verbatim ----------------
if [ -f "$HERE/ref_service_glue.c" ]; then
    pin uvm_tools.h 2b366fbdaa63bdef
    sed -n '1214,1272p;1293,1360p;1375,1597p;1656,1693p' "$UVM/uvm_gpu_replayable_faults.c" > "$TMP/x_service.h"
    sed -n '181,188p;522,537p' "$UVM/uvm_hal_types.h" > "$TMP/x_prot.h"
    sed -n '46,62p' "$UVM/uvm_tools.h" > "$TMP/x_tools_reason.h"
    gcc $CFLAGS -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_service_glue.c" "$HERE/ref_faults_blackwell.c" \
        "$HERE/ref_faults_hopper.c" "$HERE/oracle_service.c" "$HERE/oracle_thrashing.c" "$HERE/oracle_resident.c" "$HERE/oracle_prefetch.c" \
        -o "$OUT/libref_service.so"
    echo "build_ref: built $OUT/libref_service.so"
fi

# ---------------- non-replayable faults: service_managed_fault_in_block_locked, verbatim, on the service leg's shims ----------------
if [ -f "$HERE/ref_nonreplayable_glue.c" ]; then
    got="$(sha256sum "$UVM/uvm_gpu_non_replayable_faults.c" | cut -c1-16)"
    if [ "$got" != "588e72e5e3163f2e" ]; then echo "build_ref: uvm_gpu_non_replayable_faults.c sha256 prefix $got != pinned 588e72e5e3163f2e" >&2; exit 1; fi
    sed -n '345,445p' "$UVM/uvm_gpu_non_replayable_faults.c" > "$TMP/x_nonreplayable.h"
    gcc $CFLAGS -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_nonreplayable_glue.c" "$HERE/ref_faults_blackwell.c" \
        "$HERE/ref_faults_hopper.c" "$HERE/oracle_service.c" "$HERE/oracle_thrashing.c" "$HERE/oracle_resident.c" "$HERE/oracle_prefetch.c" \
        -o "$OUT/libref_nonreplayable.so"
    echo "build_ref: built $OUT/libref_nonreplayable.so"
fi

# ---------------- residency leg: uvm_va_block_make_resident and its mask algebra, verbatim ----------------
if [ -f "$HERE/ref_resident_glue.c" ]; then
    sed -n '36,80p' "$UVM/uvm_va_block_types.h" > "$TMP/x_va_block_types.h"
    sed -n '116,129p;149,159p;324,328p' "$UVM/uvm_va_block_types.h" > "$TMP/x_resident_types.h"
    sed -n '197,242p' "$UVM/uvm_va_block_types.h" > "$TMP/x_make_resident_ctx.h"
    sed -n '183,195p;584,593p;609,701p;721,734p;2406,2430p;3569,3577p;4379,4452p;4460,4536p;4538,4579p;4581,4619p;4625,4653p;4660,4847p;4877,4976p;4978,5114p;5120,5162p;5164,5332p' \
        "$UVM/uvm_va_block.c" > "$TMP/x_resident.h"
    sed -n '3099,3182p' "$UVM/uvm_va_block.c" > "$TMP/x_populate.h"
    gcc $CFLAGS -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_resident_glue.c" "$HERE/oracle_service.c" "$HERE/oracle_thrashing.c" "$HERE/oracle_resident.c" "$HERE/oracle_prefetch.c" \
        -o "$OUT/libref_resident.so"
    echo "build_ref: built $OUT/libref_resident.so"
fi

# ---------------- copy leg: the CPU page-copy loop of block_copy_pages (kmap + memcpy per page), verbatim ----------------
if [ -f "$HERE/ref_copy_glue.c" ]; then
    sed -n '3946,3991p' "$UVM/uvm_va_block.c" > "$TMP/x_copy_pages.h"
    gcc $CFLAGS -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_copy_glue.c" -o "$OUT/libref_copy.so"
    echo "build_ref: built $OUT/libref_copy.so"
fi

# ---------------- the fork's CXL P2P DMA file, whole and unmodified ----------------
P2P="$REF/src/nvidia/src/kernel/mem_mgr"
if [ -f "$HERE/ref_cxl_p2p_glue.c" ] && [ -f "$P2P/p2p_cxl.c" ]; then
    got="$(sha256sum "$P2P/p2p_cxl.c" | cut -c1-16)"
    if [ "$got" != "5526c5812ec155d1" ]; then echo "build_ref: p2p_cxl.c sha256 prefix $got != pinned 5526c5812ec155d1" >&2; exit 1; fi
    # the nine RM headers it includes become empty files (the glue defines what the file uses of them)
    for h in core/core.h gpu/gpu.h mem_mgr/p2p.h os/os.h nvport/nvport.h gpu/mem_mgr/mem_mgr.h gpu/mem_mgr/mem_desc.h \
             gpu/mem_mgr/ce_utils.h gpu/bus/kern_bus.h; do
        mkdir -p "$TMP/rmstub/$(dirname "$h")"; : > "$TMP/rmstub/$h"
    done
    gcc $CFLAGS -I"$TMP/rmstub" -I"$P2P" -I"$HERE" -I"$INC" "$HERE/ref_cxl_p2p_glue.c" -o "$OUT/libref_cxl_p2p.so"
    echo "build_ref: built $OUT/libref_cxl_p2p.so"
fi

# ---------------- PMM leg: the eviction order (pick_root_chunk_to_evict and the list moves), verbatim ----------------
if [ -f "$HERE/ref_pmm_glue.c" ]; then
    pin uvm_pmm_gpu.c 579fbaf9e0b9de07
    sed -n '1395,1414p;1416,1428p;1430,1448p;1450,1458p;1460,1500p' "$UVM/uvm_pmm_gpu.c" > "$TMP/x_pmm.h"
    gcc $CFLAGS -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_pmm_glue.c" -o "$OUT/libref_pmm.so"
    echo "build_ref: built $OUT/libref_pmm.so"
fi

# ---------------- access-counter leg: expansion + service_va_block_locked, verbatim, on the service leg's shims ----------------
if [ -f "$HERE/ref_access_counters_glue.c" ]; then
    pin uvm_gpu_access_counters.c 67b82fdfd551a698
    sed -n '198,218p;1079,1251p;1253,1260p;1287,1354p' "$UVM/uvm_gpu_access_counters.c" > "$TMP/x_access_counters.h"
    gcc $CFLAGS -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_access_counters_glue.c" "$HERE/ref_faults_blackwell.c" \
        "$HERE/ref_faults_hopper.c" "$HERE/oracle_service.c" "$HERE/oracle_thrashing.c" "$HERE/oracle_resident.c" "$HERE/oracle_prefetch.c" \
        -o "$OUT/libref_access_counters.so"
    echo "build_ref: built $OUT/libref_access_counters.so"
fi

# ---------------- thrashing detector: event callback + detection / reset, verbatim, on its own stand-ins ----------------
if [ -f "$HERE/ref_thrashing_glue.c" ]; then
    pin uvm_perf_thrashing.c 6faca5e5b498600d
    sed -n '30,43p' "$UVM/uvm_perf_utils.h" > "$TMP/x_perf_saturating.h"
    sed -n '36,119p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_thrashing_types.h"
    sed -n '376,384p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_thrashing_stamps.h"
    sed -n '1023,1073p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_thrashing_detect.h"
    sed -n '1077,1093p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_thrashing_region.h"
    sed -n '1230,1380p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_thrashing_event.h"
    gcc $CFLAGS -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_thrashing_glue.c" -o "$OUT/libref_thrashing.so"
    echo "build_ref: built $OUT/libref_thrashing.so"
fi

# ---------------- the whole thrashing heuristic (detector + hint + pin expiry), verbatim, with a persistent context ----------------
if [ -f "$HERE/ref_thrashing_hint_glue.c" ]; then
    pin uvm_perf_thrashing.c 6faca5e5b498600d
    sed -n '30,43p' "$UVM/uvm_perf_utils.h" > "$TMP/x_perf_saturating.h"
    sed -n '36,202p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_th_types.h"
    sed -n '376,394p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_th_stamps.h"
    sed -n '721,1021p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_th_throttle_pin.h"
    sed -n '1023,1073p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_th_detect.h"
    sed -n '1077,1093p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_th_region.h"
    sed -n '1172,1226p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_th_pinned_update.h"
    sed -n '1230,1380p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_th_event.h"
    sed -n '1382,1780p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_th_hint.h"
    sed -n '1822,1906p' "$UVM/uvm_perf_thrashing.c" > "$TMP/x_th_unpin_worker.h"
    sed -n '1879,1934p' "$UVM/uvm_va_space.c" > "$TMP/x_vs_closest.h"
    gcc $CFLAGS -DNDEBUG -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_thrashing_hint_glue.c" -o "$OUT/libref_thrashing_hint.so"   # NDEBUG: random streams are not coherent driver histories, the reference's sanity asserts would stop them
    echo "build_ref: built $OUT/libref_thrashing_hint.so"
fi

# ---------------- access-counter front half: parse + fetch + preprocess, verbatim, on the fault leg's shims ----------------
if [ -f "$HERE/ref_access_counters_front_glue.c" ]; then
    pin uvm_turing_access_counter_buffer.c b4afc3558d7cc578
    sed -n '481,520p' "$UVM/uvm_hal_types.h" > "$TMP/x_ac_entry.h"
    sed -n '88,198p' "$UVM/uvm_turing_access_counter_buffer.c" > "$TMP/x_ac_turing.h"
    sed -n '798,821p;834,860p;862,940p;942,969p;974,992p' "$UVM/uvm_gpu_access_counters.c" > "$TMP/x_ac_front.h"
    gcc $CFLAGS -I"$TMP" -I"$HERE" -I"$INC" -I"$UVM" "$HERE/ref_access_counters_front_glue.c" "$HERE/ref_faults_blackwell.c" \
        "$HERE/ref_faults_hopper.c" -o "$OUT/libref_access_counters_front.so"
    echo "build_ref: built $OUT/libref_access_counters_front.so"
fi
