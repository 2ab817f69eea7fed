/* C ABI of libsmpb200_core.so - host runtime of smdistributed_modelparallel_b200.
 * Sources: csrc/core/msgbus.cpp, csrc/core/runtime.cpp, csrc/symm/*.cu.  Reference call sites
 * (paths relative to smdistributed/modelparallel/) are cited where a function replaces a native
 * entry point of the reference's smplib.so. */
#ifndef SMPB200_CORE_H_
#define SMPB200_CORE_H_
#ifdef __cplusplus
extern "C" {
#endif

/* ---- message bus (smp_async_send / bcast / recv / wait / poll / retrieve: backend/collectives.py:233-324;
 *      listener: torch/server_comm.py:162-182) */
int smpc_bus_listen(void);                                              /* -> ephemeral port, every interface */
int smpc_bus_listen_on(const char* ipv4);                               /* -> ephemeral port on one interface */
int smpc_bus_set_token(unsigned long long token);                       /* job secret checked at connect */
int smpc_bus_connect(int rank, int size, const char* peers_csv, int timeout_s);
int smpc_bus_initialized(void);
int smpc_bus_rank(void);
int smpc_bus_size(void);
int smpc_bus_send(const void* buf, long long len, int dst, long long txid, int server);
int smpc_bus_bcast(const void* buf, long long len, int n, const int* ranks, long long txid, int server);
int smpc_bus_wait_sends(long long txid);                                /* txid < 0: all */
long long smpc_bus_poll(int src, long long txid);                       /* length or -1 */
long long smpc_bus_wait(int src, long long txid, int timeout_ms);       /* length, -1 timeout, -2 peer failed */
long long smpc_bus_retrieve(int src, long long txid, void* out, long long cap);
long long smpc_bus_server_peek(int* src, long long* txid, int timeout_ms);
long long smpc_bus_server_pop(void* out, long long cap);
int smpc_bus_peer_failed(void);                                         /* failing rank or -1 */
int smpc_bus_shutdown(int success);                                     /* backend/core.py:227-232 */

/* ---- gradient readiness counter (GradCounter: torch/model.py:401-403, torch/worker.py:406-417) */
void* smpc_gc_create(int num_params, int num_microbatches);
void smpc_gc_destroy(void* h);
void smpc_gc_set_microbatch(void* h, int mb);
int smpc_gc_get_microbatch(void* h);
int smpc_gc_increment_expected(void* h, int mb, const int* ids, int n);
int smpc_gc_mark_fwd_done(void* h, int mb);
int smpc_gc_mark_grad_computed(void* h, int id);                        /* 1 = gradient final for the step */
int smpc_gc_is_ready(void* h, int id);
int smpc_gc_expected(void* h, int mb, int id);
int smpc_gc_seen(void* h, int mb, int id);
void smpc_gc_expected_totals(void* h, int* out);
void smpc_gc_clear(void* h);

/* ---- bucket planner (compute_bucket_assignment_by_size: torch/ddp_model.py:336-341) */
int smpc_plan_buckets(int n, const long long* nbytes, const int* dtype_key, long long first_cap, long long cap,
                      int* bucket_of);

/* ---- timeline + metrics (backend/core.py:290,524-562) */
int smpc_timeline_create(const char* path, int rank);
int smpc_timeline_enabled(void);
void smpc_timeline_start_step(void);
void smpc_timeline_end_step(void);
void smpc_timeline_record_pipeline_event(int microbatch, const char* label);
int smpc_timeline_num_events(void);
int smpc_timeline_flush(void);
void smpc_timeline_destroy(void);
void smpc_pool_record(long long delta_alloc, long long delta_reserved, int ok);
int smpc_get_and_reset_memory_metrics(double* out4);
int smpc_get_and_reset_alloc_metrics(long long* out2);
int smpc_device_count(void);
/* in-library self-test of the counter, the bucket planner and the metrics (the reference's library carries
 * native self-tests too: test/backend/d2d_test.py:11, from_blob_tensor_test.py:11); 0 = ok */
int smpc_selftest_runtime(char* msg, int cap);

/* ---- symmetric (peer-addressable) memory over CUDA IPC (no counterpart among the reference's Python-visible
 *      entry points; replaces NCCL for tensor-parallel traffic: torch/nn/utils.py:701-821) */
const char* smpc_symm_last_error(void);
int smpc_symm_alloc(long long bytes, void** dev_ptr, unsigned char* handle64_out);
int smpc_symm_open(const unsigned char* handle64, void** peer_ptr);
int smpc_symm_close(void* peer_ptr);
int smpc_symm_free(void* dev_ptr);
int smpc_symm_can_access_peer(int peer_device);
int smpc_symm_selftest(long long bytes);                                /* needs a GPU; 0 = ok */

#ifdef __cplusplus
}
#endif
#endif /* SMPB200_CORE_H_ */
