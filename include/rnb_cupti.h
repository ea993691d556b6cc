/* rnb_cupti -- kernel timeline of the current process through the CUPTI Activity API.
 *
 * C ABI of librnb_cupti.so, the B200-era counterpart of the reference's only native component,
 * utils/cupti.cpp (a pybind11 module with initialize / flush / report, reference cupti.cpp:142-166,
 * built by build_cupti.sh:19-23).  Kept in a library of its own so that librnb_b200.so does not
 * depend on libcupti.  Differences from the reference, all deliberate:
 *   - records are read as CUpti_ActivityKernel9 (CUDA 12.9) instead of assuming the 144-byte
 *     CUpti_ActivityKernel4 of CUDA 9 (reference cupti.cpp:19,73);
 *   - CUPTI_ACTIVITY_KIND_CONCURRENT_KERNEL is enabled, not KIND_KERNEL (cupti.cpp:145), which
 *     serialises kernel execution and would hide the programmatic-dependent-launch overlap;
 *   - failures come back as codes + rnb_cupti_last_error(), the reference calls exit(-1)
 *     (cupti.cpp:25-49); the record list is guarded by a mutex (CUPTI delivers buffers on its own
 *     thread).
 * All functions return 0 on success and a negative value on failure.
 */
#ifndef RNB_CUPTI_H_
#define RNB_CUPTI_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Start collecting one record per kernel launched by this process from now on.
 * Replaces cupti.initialize() (reference utils/cupti.cpp:142-151). */
int rnb_cupti_initialize(void);

/* Make CUPTI hand over the records of every kernel that has completed so far.  Call after
 * synchronising the streams of interest.  Replaces cupti.flush() (utils/cupti.cpp:154-156). */
int rnb_cupti_flush(void);

/* Number of records collected and not yet taken: the length of the list cupti.report() would
 * return (utils/cupti.cpp:160-166). */
int rnb_cupti_count(void);

/* Copy record `index` (0 <= index < rnb_cupti_count()): the kernel's (demangled when possible) name
 * into name[0..name_cap), start / end in nanoseconds on CUPTI's clock, and the launch geometry.
 * Any output pointer may be NULL.  One element of the list cupti.report() returns
 * (utils/cupti.cpp:160-166); the geometry fields have no counterpart in the reference. */
int rnb_cupti_get(int index, char* name, int name_cap, uint64_t* start_ns, uint64_t* end_ns,
                  uint32_t* device, uint32_t* stream, int32_t grid3[3], int32_t block3[3]);

/* Drop the collected records.  rnb_cupti_get + rnb_cupti_clear together replace cupti.report()
 * (utils/cupti.cpp:160-166), which returns the list and empties it. */
int rnb_cupti_clear(void);

/* Stop collecting (no counterpart in the reference, which never disables the activity kind). */
int rnb_cupti_finalize(void);

const char* rnb_cupti_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* RNB_CUPTI_H_ */
