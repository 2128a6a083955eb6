/* libt2b200 -- C ABI of the B200-native Tacotron2 acoustic-model hot path.
 *
 * The reference (mattm458/tacotron2) is pure Python/PyTorch and has no FFI of its own; the
 * boundary this library sits behind is the reference's nn.Module API.  Every entry point
 * below names the reference function it replaces (paths relative to the reference root);
 * tacotron2_b200/model.py mirrors those modules and binds these symbols with ctypes
 * (INTEGRATION.md shows the stub).
 *
 * Conventions: all pointers are DEVICE pointers unless stated; tensors are dense, row-major,
 * fp32 unless stated; `stream` is a cudaStream_t passed as void*; calls are asynchronous on
 * that stream and never synchronise the host; no allocation or ownership transfer happens
 * inside a call (outputs and a workspace of t2_workspace_bytes() are the caller's); the
 * return value is 0 on success or a negative T2_ERR_* code, with t2_last_error() giving a
 * thread-local message.  A context is not thread-safe.  There is no CPU fallback.
 */
#ifndef T2B200_H
#define T2B200_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct t2_ctx t2_ctx;

/* Constructor arguments of Tacotron2 (model/tacotron2.py:15-34). */
typedef struct {
  int num_chars, encoded_dim, encoder_kernel_size, num_mels, prenet_dim;
  int att_rnn_dim, att_dim, rnn_hidden_dim, postnet_dim;
  float dropout;
  int num_speakers;     /* 0 = speaker tokens off */
  int controls_dim;     /* 0 = controls off */
  int description_dim;  /* 0 = description embeddings off (else 768 -> Linear -> 128) */
  int precision;        /* gate GEMMs: 0 = fp32-parity (three bf16 hi/lo passes, fp32 accumulate), 1 = bf16 */
} t2_model_config;

const char* t2_last_error(void);
int t2_abi_version(void);

/* Context = one model replica on the current device. */
int t2_ctx_create(const t2_model_config* cfg, t2_ctx** out);
int t2_ctx_destroy(t2_ctx* ctx);
/* Register the device pointer of one fp32 parameter / buffer under its reference state-dict
 * key ("decoder.att_rnn.weight_ih", ...).  Zero-copy: the pointer must stay valid. */
int t2_ctx_set_param(t2_ctx* ctx, const char* key, const float* ptr, long long numel);
/* (Re)build the packed copies the kernels read (call after loading or updating weights). */
int t2_ctx_commit(t2_ctx* ctx, void* stream);
/* Upper bound of the workspace any call below needs for these sizes. */
long long t2_workspace_bytes(const t2_ctx* ctx, int B, int L, int T);

/* Encoder.forward (model/encoder.py:54-67).  chars_idx (B,L) int64, lens (B) int64 ->
 * encoded (B,L,E).  keep: NULL (eval) or train-mode dropout keep-masks (3,B,L,E) uint8.
 * train must be 0 here (eval: BatchNorm running statistics folded into the conv epilogue); the train-mode forward with
 * batch statistics and a tape for the backward is t2_encoder_train_fwd below. */
int t2_encoder_fwd(t2_ctx* ctx, const long long* chars_idx, const long long* lens, int B, int L, int train,
                   const unsigned char* keep, float* encoded, void* ws, long long ws_bytes, void* stream);

/* Conditioning + attention memory projection (model/tacotron2.py:201-212,229).
 * encoded (B,L,E), speaker_id (B) int64 or NULL, description (B,Ddesc) or NULL ->
 * memory (B,L,EF), processed_memory (B,L,A). */
int t2_memory_fwd(t2_ctx* ctx, const float* encoded, const long long* speaker_id, const float* description, int B, int L,
                  float* memory, float* processed_memory, void* ws, long long ws_bytes, void* stream);

/* The autoregressive frame loop of Tacotron2.forward with teacher_forcing=False
 * (model/tacotron2.py:262-325) as ONE persistent cooperative kernel: prenet, attention
 * LSTMCell, location-sensitive attention, decoder LSTMCell, mel/gate projection, stop logic.
 * mask (B,L) uint8 1 = padded; controls (B,C) or NULL; prenet_keep (max_len+1,2,B,P) uint8
 * keep-masks for AlwaysDropout or NULL (in-kernel RNG from `seed`).
 * Outputs: mels (B,max_len,M), gates (B,max_len), alignments (B,max_len,L) -- rows >= *n_frames
 * are left untouched; lengths (B) int64; n_frames (1) int32 (device). B <= 64 per call. */
int t2_decode_ar(t2_ctx* ctx, const float* memory, const float* processed_memory, const unsigned char* mask,
                 const float* controls, int B, int L, int max_len, const unsigned char* prenet_keep,
                 unsigned long long seed, float* mels, float* gates, float* alignments, long long* lengths,
                 int* n_frames, void* ws, long long ws_bytes, void* stream);

/* One chunk (<= 64 utterances) of a larger autoregressive batch.  The reference's stop condition couples the whole batch: the loop
 * ends when EVERY utterance is done (model/tacotron2.py:321-322), finished utterances keep decoding and `lengths` keeps counting
 * (:319-320).  A batch of more than 64 utterances is decoded chunk by chunk; a chunk whose own utterances finish early must still
 * produce as many frames as the slowest chunk, so the stop condition is honoured only once `min_frames` frames have been emitted.
 * min_frames = 0 is t2_decode_ar. */
int t2_decode_ar_chunk(t2_ctx* ctx, const float* memory, const float* processed_memory, const unsigned char* mask,
                       const float* controls, int B, int L, int max_len, int min_frames, const unsigned char* prenet_keep,
                       unsigned long long seed, float* mels, float* gates, float* alignments, long long* lengths,
                       int* n_frames, void* ws, long long ws_bytes, void* stream);

/* Teacher-forced frame loop (model/tacotron2.py:255-258,276-317): prenet over the shifted target
 * in bulk, then the same persistent kernel without prenet/stop phases.  mel_target (B,T,M);
 * train != 0 applies the p=0.1 hidden-state dropouts (att_keep/rnn_keep (T,B,H) or RNG). */
int t2_decode_tf(t2_ctx* ctx, const float* memory, const float* processed_memory, const unsigned char* mask,
                 const float* controls, const float* mel_target, int B, int L, int T, const unsigned char* prenet_keep,
                 int train, const unsigned char* att_keep, const unsigned char* rnn_keep, unsigned long long seed,
                 float* mels, float* gates, float* alignments, void* ws, long long ws_bytes, void* stream);

/* Decoder.forward, one frame from explicit state (model/decoder.py:53-119).  All state tensors
 * are (B,dim) row-major; w / w_cum (B,L); *_out may alias nothing.  w_cum_out = w_cum + w_out. */
int t2_decoder_step(t2_ctx* ctx, const float* prev_mel_prenet, const float* att_h, const float* att_c, const float* att_ctx,
                    const float* w, const float* w_cum, const float* rnn_h, const float* rnn_c, const float* memory,
                    const float* processed_memory, const unsigned char* mask, const float* controls, int B, int L,
                    float* mel_out, float* gate_out, float* att_h_out, float* att_c_out, float* att_ctx_out, float* w_out,
                    float* w_cum_out, float* rnn_h_out, float* rnn_c_out, void* ws, long long ws_bytes, void* stream);

/* Attention.forward, stand-alone (model/attention.py:34-69): attention_hidden_state (B,H), memory (B,L,EF),
 * processed_memory (B,L,A), attention_weights_cat (B,2,L) [channel 0 = previous weights, 1 = cumulative, model/decoder.py:78-80],
 * mask (B,L) uint8 1 = padded -> context (B,EF), weights (B,L).  The same arithmetic runs fused inside the persistent decoder
 * kernel; this entry exists so the module boundary is callable on its own.  No workspace. */
int t2_attention_step(t2_ctx* ctx, const float* attention_hidden_state, const float* memory, const float* processed_memory,
                      const float* attention_weights_cat, const unsigned char* mask, int B, int L, float* context, float* weights,
                      void* stream);

/* prenet (model/tacotron2.py:85-92) on N rows: x (N,M) -> out (N,P); keep1/keep2 (N,P) uint8 or NULL (RNG). */
int t2_prenet_fwd(t2_ctx* ctx, const float* x, const unsigned char* keep1, const unsigned char* keep2,
                  unsigned long long seed, int N, float* out, void* ws, long long ws_bytes, void* stream);

/* Postnet.forward (model/postnet.py:51-52) on channels-last input: mels (B,T,M) with row stride
 * t_stride frames per batch item -> out (B,T,M) dense.  keep: NULL or (5,B,T,postnet_dim) uint8. */
int t2_postnet_fwd(t2_ctx* ctx, const float* mels, int t_stride, int B, int T, int train, const unsigned char* keep,
                   float* out, void* ws, long long ws_bytes, void* stream);

/* Residual + output masking (model/tacotron2.py:333-345), in place: post <- masked(mels + post),
 * mels <- masked(mels), gates <- masked_fill(-1000).  lengths (B) int64. */
int t2_finalize(t2_ctx* ctx, float* mels, float* post, float* gates, const long long* lengths, int B, int T, int t_stride,
                void* stream);

/* ------------------------------------------------------------------------------------------------------------------
 * Training path.  The reference gets its backward pass from autograd over
 * Tacotron2.forward (training_step, model/tts_model.py:165-202); here every module has a train-mode forward that records
 * a caller-owned "tape" (sized by t2_*_tape_bytes) and a backward that consumes it.  Parameter gradients are WRITTEN
 * (not accumulated) to the destinations registered per state-dict key with t2_ctx_set_grad.  Dropout keep-masks follow the
 * forward's convention: explicit uint8 masks or NULL = counter RNG from `seed` (the backward regenerates the same draws).
 * ------------------------------------------------------------------------------------------------------------------ */
int t2_ctx_set_grad(t2_ctx* ctx, const char* key, float* ptr, long long numel);   /* ptr NULL unregisters */
long long t2_train_workspace_bytes(const t2_ctx* ctx, int B, int L, int T);

/* Postnet.forward under train() (model/postnet.py:8-52): BatchNorm batch statistics over all (b,t), running statistics
 * updated in place (momentum 0.1), dropout after every layer.  keep (5,B,T,postnet_dim) or NULL. */
long long t2_postnet_tape_bytes(const t2_ctx* ctx, int B, int T);
int t2_postnet_train_fwd(t2_ctx* ctx, const float* mels, int t_stride, int B, int T, const unsigned char* keep, unsigned long long seed,
                         float* out, void* tape, long long tape_bytes, void* ws, long long ws_bytes, void* stream);
/* d_out (B,T,M): gradient w.r.t. the postnet output; d_mels_in (B,T,M) or NULL is added to the input gradient;
 * d_mels (B,T,M) receives the total.  Writes postnet.postnet.* gradients. */
int t2_postnet_bwd(t2_ctx* ctx, int B, int T, const unsigned char* keep, unsigned long long seed, const float* d_out,
                   const float* d_mels_in, const void* tape, long long tape_bytes, float* d_mels, void* ws, long long ws_bytes,
                   void* stream);

/* Encoder.forward under train() (model/encoder.py:54-67) and its backward.  keep (3,B,L,E) uint8 or NULL (counter RNG).
 * `encoded` in t2_encoder_bwd is the forward's output (the BiLSTM hidden states, needed for the recurrent weight gradient). */
long long t2_encoder_tape_bytes(const t2_ctx* ctx, int B, int L);
int t2_encoder_train_fwd(t2_ctx* ctx, const long long* chars_idx, const long long* lens, int B, int L, const unsigned char* keep,
                         unsigned long long seed, float* encoded, void* tape, long long tape_bytes, void* ws, long long ws_bytes,
                         void* stream);
int t2_encoder_bwd(t2_ctx* ctx, const long long* chars_idx, const long long* lens, int B, int L, const unsigned char* keep,
                   unsigned long long seed, const float* encoded, const float* d_encoded, const void* tape, long long tape_bytes, void* ws,
                   long long ws_bytes, void* stream);
/* Backward of t2_memory_fwd (conditioning + att_encoder, model/tacotron2.py:201-212,229): d_memory (B,L,EF),
 * d_processed_memory (B,L,A) -> d_encoded (B,L,E); writes att_encoder / speaker_embedding / description_embeddings_linear
 * gradients.  `memory` is the forward's output. */
int t2_memory_bwd(t2_ctx* ctx, const float* memory, const long long* speaker_id, const float* description, const float* d_memory,
                  const float* d_processed_memory, int B, int L, float* d_encoded, void* ws, long long ws_bytes, void* stream);

/* Teacher-forced frame loop under train() (model/tacotron2.py:255-258,276-317; hidden-state dropouts p = 0.1 active,
 * model/decoder.py:75,101): same persistent kernel in history mode.  Arguments as t2_decode_tf; tape from t2_decode_tape_bytes. */
long long t2_decode_tape_bytes(const t2_ctx* ctx, int B, int L, int T);
int t2_decode_tf_train_fwd(t2_ctx* ctx, const float* memory, const float* processed_memory, const unsigned char* mask,
                           const float* controls, const float* mel_target, int B, int L, int T, const unsigned char* prenet_keep,
                           const unsigned char* att_keep, const unsigned char* rnn_keep, unsigned long long seed, float* mels,
                           float* gates, float* alignments, void* tape, long long tape_bytes, void* ws, long long ws_bytes, void* stream);
/* Backward (BPTT) of the frame loop: one persistent kernel for the recurrent chain + bulk tensor-core GEMMs for everything
 * off the chain.  d_mels (B,T,M), d_gates (B,T): gradients w.r.t. the raw decoder outputs; alignments (B,T,L) from the forward.
 * Outputs d_memory (B,L,EF), d_processed_memory (B,L,A); writes decoder.* and prenet.* parameter gradients. */
int t2_decode_tf_bwd(t2_ctx* ctx, const float* memory, const float* processed_memory, const unsigned char* mask, const float* mel_target,
                     int B, int L, int T, const unsigned char* att_keep, const unsigned char* rnn_keep, unsigned long long seed,
                     const float* alignments, const float* d_mels, const float* d_gates, const void* tape, long long tape_bytes,
                     float* d_memory, float* d_processed_memory, void* ws, long long ws_bytes, void* stream);

/* training_step loss (model/tts_model.py:197-201): BCEWithLogits(gates) + MSE(mels) + MSE(mels_post), plain means over the
 * padded tensors; loss (1) device scalar; d_mels/d_post (B,T,M), d_gates (B,T) may be NULL.  ws >= 4 KB. */
int t2_loss_fwd_bwd(t2_ctx* ctx, const float* mels, const float* post, const float* gates, const float* mel_target,
                    const float* gate_target, int B, int T, float* loss, float* d_mels, float* d_post, float* d_gates, void* ws,
                    long long ws_bytes, void* stream);
/* Backward of t2_finalize (model/tacotron2.py:333-345).  Inputs may be NULL (= zero gradient). */
int t2_finalize_bwd(t2_ctx* ctx, const float* d_mels, const float* d_post, const float* d_gates, const long long* lengths, int B, int T,
                    float* d_mels_raw, float* d_postnet_out, float* d_gates_raw, void* stream);

/* Training-step shell: torch.nn.utils.clip_grad_norm_(params, max_norm) followed by
 * torch.optim.Adam(lr, betas, eps, weight_decay).step() -- the optimizer contract of model/tts_model.py:78-91 (Adam, lr 1e-3,
 * L2 weight decay 1e-6) with run/train.py:239-240 (gradient_clip_val 1.0) -- over a list of fp32 device tensors, three launches,
 * no host synchronisation (the clip coefficient stays on the device).  params / grads / exp_avg / exp_avg_sq: host arrays of
 * n_tensors device pointers; numel: host array.  step = 1 for the first update (bias correction).  max_norm <= 0 disables
 * clipping.  table_cached != 0: the tensor count and sizes are the same as in the previous call on this workspace (skips the
 * re-upload of the chunk table).  grad_norm_out: device float or NULL, receives the pre-clip global 2-norm.
 * ws: t2_clip_adam_workspace_bytes(n_tensors, t2_clip_adam_chunks(n_tensors, numel)) bytes, kept by the caller across steps. */
long long t2_clip_adam_chunks(int n_tensors, const long long* numel);
long long t2_clip_adam_workspace_bytes(int n_tensors, long long total_chunks);
int t2_clip_adam_step(int n_tensors, float* const* params, float* const* grads, float* const* exp_avg, float* const* exp_avg_sq,
                      const long long* numel, float lr, float beta1, float beta2, float eps, float weight_decay, float max_norm,
                      long long step, int table_cached, float* grad_norm_out, void* ws, long long ws_bytes, void* stream);

/* Batched log-mel extraction: TacotronMelSpectrogram(n_mels, sample_rate)(wav) of datasets/tts_dataset.py:166-168,204
 * (implementation external to the reference: speech_utils, un-vendored; parameters as stated in
 * datasets/prosody_dataset.py:39-50,66-67 and run/say.py:161-171: n_fft 1024, win 1024, hop 256, center + reflect pad,
 * periodic Hann, magnitude, slaney filterbank, log(clamp(1e-5)), transposed to (T, n_mels)).
 * wav (B, wav_stride) fp32 zero padded, wav_len (B) int64 samples; window (n_fft); fb (n_mels, n_fft/2+1) fp32 filterbank,
 * fb_range (n_mels, 2) int32 [lo,hi) non-zero bin range per filter.  out (B, t_stride, n_mels): frames
 * t < 1 + wav_len/hop hold log-mels, the rest zeros (pad_sequence, datasets/tts_dataloader.py:28);
 * n_frames (B) int64 or NULL receives the frame counts (mel_spectrogram_len, datasets/tts_dataset.py:209). */
int t2_melspec(const float* wav, const long long* wav_len, int B, long long wav_stride, const float* window, const float* fb,
               const int* fb_range, int n_fft, int hop, int n_mels, float* out, int t_stride, long long* n_frames, void* stream);

/* Silence trim bounds of a waveform batch: librosa.effects.trim(wav, top_db, frame_length) as datasets/tts_dataset.py:194-200 calls it
 * (hop_length 512 is librosa's default): start / end (B) int64 sample bounds of the kept interval of every utterance (end = start = 0
 * when no frame is above the threshold).  wav (B, wav_stride) fp32, wav_len (B) int64 on the device, max_len >= every length.
 * librosa is external to the reference: parity is pinned to oracle/trim_oracle.py.  ws: t2_trim_workspace_bytes. */
long long t2_trim_workspace_bytes(int B, long long max_len, int hop_length);
int t2_trim(const float* wav, const long long* wav_len, int B, long long wav_stride, long long max_len, float top_db, int frame_length,
            int hop_length, long long* start, long long* end, void* ws, long long ws_bytes, void* stream);

/* t2_trim without a workspace argument and without any allocation, for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  This entry point has no counterpart in the reference.
 * Same inputs, outputs and results as t2_trim; the frame-RMS scratch is a file-scope device array
 * of T2_TRIM_STATIC_MAX_FRAMES floats (1 MiB, resident on every GPU that loads the library until the process exits).  Requests with
 * B * (1 + max_len / hop_length) > T2_TRIM_STATIC_MAX_FRAMES return T2_ERR_WORKSPACE (-4) with a message in t2_last_error() and
 * launch nothing.  All calls on a device share the array (it keeps the previous call's contents; nothing in it needs clearing):
 * calls that overlap on the GPU race -- serialise them on one stream, or use t2_trim with a workspace per stream.  The first call
 * on each device looks the array's device address up (cudaGetSymbolAddress, no allocation). */
#define T2_TRIM_STATIC_MAX_FRAMES 262144
int t2_trim_static(const float* wav, const long long* wav_len, int B, long long wav_stride, long long max_len, float top_db,
                   int frame_length, int hop_length, long long* start, long long* end, void* stream);

/* Batch assembly: `_collate` of datasets/tts_dataloader.py:8-35 -- nn.utils.rnn.pad_sequence(batch_first=True) over every data
 * key of the batch (chars_idx (L,) int64, mel_spectrogram (T, n_mels) fp32, gate (T, 1) fp32, ...) -- plus the gate targets of
 * datasets/tts_dataset.py:211-214 (ones, last valid frame 0), in one launch.  Field f: src[f] = the B items' rows concatenated
 * (device memory, item b = rows off[f][b] .. off[f][b+1]), row_bytes[f] bytes per row (multiple of 4; any dtype),
 * out[f] (B, max_rows[f], row_bytes[f]): rows beyond an item's length are zero.  The four tables are HOST arrays of n_fields
 * entries (n_fields <= 8); off[f] points to (B+1) int64 on the device.  gate (B, gate_T) fp32 with gate_len (B) int64 on the
 * device, or both NULL: gate[b][t] = 1 for t < gate_len[b]-1, else 0.  Asynchronous on `stream`; no allocation. */
int t2_collate(int n_fields, const void* const* src, const long long* const* off, const long long* row_bytes,
               const long long* max_rows, void* const* out, int B, const long long* gate_len, float* gate, long long gate_T,
               void* stream);

/* ------------------------------------------------------------------------------------------------------------------
 * HiFi-GAN generator: model/hifi_gan.py:154-216 Generator.forward with ResBlock1 / ResBlock2
 * (:20-151), the vocoder run/say.py:153-157 and run/test.py:170 call on the acoustic model's output.  mel (B, num_mels, T) fp32
 * (the layout the reference passes: mel_spectrogram_post.swapaxes(1, 2)) -> wav (B, T * prod(upsample_rates)) fp32 = the
 * reference's (B, 1, T') output without the singleton channel.  Parameters are registered under the reference's state-dict keys
 * AFTER Generator.remove_weight_norm() (run/say.py:88): "conv_pre.weight" (C0, num_mels, 7), "conv_pre.bias", "ups.i.weight"
 * (Cin, Cout, k) [ConvTranspose1d layout], "ups.i.bias", "resblocks.n.convs1.m.{weight,bias}", "resblocks.n.convs2.m.*"
 * (ResBlock2: "resblocks.n.convs.m.*"), "conv_post.weight" (1, C, 7), "conv_post.bias".  Same conventions as above: device
 * pointers, zero-copy parameters, caller-owned workspace, asynchronous on `stream`.
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct t2_hifigan t2_hifigan;
typedef struct {
  int num_mels;                     /* 80 (conv_pre input channels, model/hifi_gan.py:161) */
  int upsample_initial_channel;     /* h.upsample_initial_channel */
  int num_upsamples;                /* len(h.upsample_rates) <= 8 */
  int upsample_rates[8];
  int upsample_kernel_sizes[8];
  int resblock;                     /* 1 = ResBlock1 (h.resblock == "1"), 2 = ResBlock2 */
  int num_kernels;                  /* len(h.resblock_kernel_sizes) <= 8 */
  int resblock_kernel_sizes[8];
  int resblock_dilation_sizes[8][3];
  int precision;                    /* 0 = fp32-parity (three bf16 hi/lo passes), 1 = bf16 */
} t2_hifigan_config;
int t2_hifigan_create(const t2_hifigan_config* cfg, t2_hifigan** out);
int t2_hifigan_destroy(t2_hifigan* g);
int t2_hifigan_set_param(t2_hifigan* g, const char* key, const float* ptr, long long numel);
int t2_hifigan_commit(t2_hifigan* g, void* stream);
long long t2_hifigan_workspace_bytes(const t2_hifigan* g, int B, int T);
long long t2_hifigan_out_len(const t2_hifigan* g, int T);
int t2_hifigan_fwd(t2_hifigan* g, const float* mel, int B, int T, float* wav, void* ws, long long ws_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------------------
 * Description-embedding front end: BertModel(input_ids, token_type_ids, attention_mask).pooler_output
 * as run/say.py:93-106 computes it (google-bert/bert-base-uncased), batched: input_ids / token_type_ids (B, S) int64 (token_type_ids
 * may be NULL = zeros), attention_mask (B, S) fp32 1 = token / 0 = padding (NULL = all ones) -> pooler_output (B, hidden) and
 * optionally last_hidden_state (B, S, hidden).  Parameters under the Hugging Face BertModel state-dict keys
 * ("embeddings.word_embeddings.weight", "encoder.layer.0.attention.self.query.weight", ..., "pooler.dense.bias").  The WordPiece
 * tokenizer is host string processing outside this library.
 * ------------------------------------------------------------------------------------------------------------------ */
typedef struct t2_bert t2_bert;
typedef struct {
  int vocab_size, hidden_size, num_layers, num_heads, intermediate_size, max_position, type_vocab_size;
  float layer_norm_eps;
  int precision;   /* 0 = fp32-parity (three bf16 hi/lo passes), 1 = bf16 */
} t2_bert_config;
int t2_bert_create(const t2_bert_config* cfg, t2_bert** out);
int t2_bert_destroy(t2_bert* g);
int t2_bert_set_param(t2_bert* g, const char* key, const float* ptr, long long numel);
int t2_bert_commit(t2_bert* g, void* stream);
long long t2_bert_workspace_bytes(const t2_bert* g, int B, int S);
int t2_bert_pooler_fwd(t2_bert* g, const long long* input_ids, const long long* token_type_ids, const float* attention_mask, int B, int S,
                       float* pooler_output, float* last_hidden_state, void* ws, long long ws_bytes, void* stream);

/* Sticky device status of the context's persistent kernels (decoder forward / backward).  They bound every wait; a timeout makes
 * all CTAs leave together, NaN-fills the launch's outputs and sets this word.  *status_out = 0 or T2_ERR_DEVICE_ABORT; reading
 * clears it.  Synchronises the device: call it where the host synchronises anyway (reading a loss, n_frames, a checkpoint). */
int t2_ctx_status(t2_ctx* ctx, int* status_out);

/* Debug: CUDA events around the persistent decoder kernels on the launching stream.  which: 0 = decoder_persistent_kernel,
 * 1 = decoder_bwd_kernel of the last timed call; t2_debug_kernel_ms synchronises on the end event. */
int t2_debug_set_timing(t2_ctx* ctx, int on);
int t2_debug_kernel_ms(t2_ctx* ctx, int which, float* ms_out);

/* Debug: SM-clock cycles CTA 0 spent in each phase of the last decode call (host pointer, 16 slots;
 * synchronises the device).  Slots: 0 att-LSTM tail, 1 query, 2 attention tail, 3 dec-LSTM tail, 4 projection, 5 prenet1,
 * 6 prenet2, 7 extra barrier (T2B200_DEBUG_BARRIER), 8/9 att/dec gate GEMM, 10-14 attention conv/dense/sync/softmax/context. */
int t2_debug_phase_cycles(t2_ctx* ctx, unsigned long long* host_out16);

/* Debug: out (rows,64) = W (rows,K) * x (64,K)^T through the decoder's tcgen05 split-K path (TMA ring,
 * cluster of 4, DSMEM reduce).  npass 3 = fp32-parity bf16 hi/lo split, 1 = plain bf16.  Synchronous. */
int t2_debug_tcgemm(const float* W, const float* x, float* out, int rows, int K, int npass, void* ws, long long ws_bytes,
                    void* stream);

#ifdef __cplusplus
}
#endif
#endif
