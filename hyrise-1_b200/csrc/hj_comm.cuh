// hj_comm.cuh — multi-GPU exchange. Placeholder until the NCCL all-to-all lands: the entry points exist
// (include/hj.h declares them) and fail loudly.
#pragma once

extern "C" {

int hj_comm_unique_id(uint8_t id[HJ_NCCL_UNIQUE_ID_BYTES]) {
  (void)id;
  return HJ_ERR_COMM;
}

int hj_comm_init(hj_ctx* ctx, const uint8_t id[HJ_NCCL_UNIQUE_ID_BYTES], int n_ranks, int rank) {
  (void)id; (void)n_ranks; (void)rank;
  return fail(ctx, HJ_ERR_COMM, "multi-GPU exchange not built yet");
}

int hj_join_sharded(hj_ctx* ctx, const hj_relation* build_shard, const hj_relation* probe_shard, int mode,
                    void* cuda_stream, hj_result_info* out_info) {
  (void)build_shard; (void)probe_shard; (void)mode; (void)cuda_stream; (void)out_info;
  return fail(ctx, HJ_ERR_COMM, "multi-GPU exchange not built yet");
}

}  // extern "C"
