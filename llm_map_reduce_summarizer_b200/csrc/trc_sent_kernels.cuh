// Sentence path kernels + host staging (placeholder until the Punkt classifier lands).
static int sentence_stage_count(trc_ctx* ctx, cudaStream_t, int32_t*, int32_t nraw) {
  ctx->err = "over-budget un-merged segment: sentence path not built yet (" + std::to_string(nraw) + " segments)";
  return TRC_EUNSUPPORTED;
}
static int sentence_stage_write(trc_ctx*, cudaStream_t, const int32_t*) { return TRC_OK; }
static int sentence_stage_emit(trc_ctx*, cudaStream_t) { return TRC_OK; }
