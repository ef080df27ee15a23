/* A serving-runtime style caller of the C ABI, in plain C (no Python, no torch): what
 * runtime/core/decoder/torch_asr_model.cc:300-319 (TorchAsrModel::ForwardEncoderFunc ->
 * TorchScript `forward_greedy_search`) would call instead of the scripted greedy search.
 *
 *   greedy_caller <fixture.bin> <out.bin>
 *
 * fixture.bin (little endian, written by tests/test_abi_c_gpu.py):
 *   int32  B T V Emb H L J blank n_steps max_tokens
 *   float  enc_proj[B*T*J]; int32 lens[B]; float embed[V*Emb];
 *   per layer l < L: float w_ih[4H*(l?H:Emb)], w_hh[4H*H], b_ih[4H], b_hh[4H];
 *   float  w_c[J*H], b_c[J], w_out[V*J], b_out[V]
 * out.bin: int32 rounds, int32 n_tok[B], int32 tokens[B*max_tokens]
 *
 * Built by the test with: gcc -I include -I $CUDA/include greedy_caller.c libb200_transducer.so -lcudart
 */
#include <cuda_runtime_api.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "b200_transducer.h"

static void* dev_from_file(FILE* f, size_t bytes) {
  void* h = malloc(bytes);
  void* d = NULL;
  if (!h || fread(h, 1, bytes, f) != bytes) {
    fprintf(stderr, "short read (%zu bytes)\n", bytes);
    exit(2);
  }
  if (cudaMalloc(&d, bytes) != cudaSuccess || cudaMemcpy(d, h, bytes, cudaMemcpyHostToDevice) != cudaSuccess) {
    fprintf(stderr, "cudaMalloc/cudaMemcpy failed\n");
    exit(3);
  }
  free(h);
  return d;
}

int main(int argc, char** argv) {
  if (argc != 3) {
    fprintf(stderr, "usage: %s fixture.bin out.bin\n", argv[0]);
    return 1;
  }
  FILE* f = fopen(argv[1], "rb");
  if (!f) return 1;
  int32_t hd[10];
  if (fread(hd, 4, 10, f) != 10) return 2;
  const int B = hd[0], T = hd[1], V = hd[2], Emb = hd[3], H = hd[4], L = hd[5], J = hd[6], blank = hd[7],
            n_steps = hd[8], max_tokens = hd[9];
  if (b200_compiled_arch() != 100) {
    fprintf(stderr, "library not built for sm_100a\n");
    return 4;
  }
  const float* enc_proj = (const float*)dev_from_file(f, sizeof(float) * (size_t)B * T * J);
  const int32_t* lens = (const int32_t*)dev_from_file(f, sizeof(int32_t) * (size_t)B);
  const float* embed = (const float*)dev_from_file(f, sizeof(float) * (size_t)V * Emb);
  const float** lstm = (const float**)malloc(sizeof(float*) * 4 * (size_t)L);
  for (int l = 0; l < L; ++l) {
    const size_t in = l ? (size_t)H : (size_t)Emb;
    lstm[4 * l + 0] = (const float*)dev_from_file(f, sizeof(float) * 4 * H * in);
    lstm[4 * l + 1] = (const float*)dev_from_file(f, sizeof(float) * 4 * (size_t)H * H);
    lstm[4 * l + 2] = (const float*)dev_from_file(f, sizeof(float) * 4 * (size_t)H);
    lstm[4 * l + 3] = (const float*)dev_from_file(f, sizeof(float) * 4 * (size_t)H);
  }
  const float* w_c = (const float*)dev_from_file(f, sizeof(float) * (size_t)J * H);
  const float* b_c = (const float*)dev_from_file(f, sizeof(float) * (size_t)J);
  const float* w_out = (const float*)dev_from_file(f, sizeof(float) * (size_t)V * J);
  const float* b_out = (const float*)dev_from_file(f, sizeof(float) * (size_t)V);
  fclose(f);

  int32_t *tokens = NULL, *n_tok = NULL;
  void* ws = NULL;
  const size_t ws_bytes = b200_greedy_workspace_bytes(B, V, H, L, J);
  cudaMalloc((void**)&tokens, sizeof(int32_t) * (size_t)B * max_tokens);
  cudaMalloc((void**)&n_tok, sizeof(int32_t) * (size_t)B);
  cudaMemset(n_tok, 0, sizeof(int32_t) * (size_t)B);
  cudaMalloc(&ws, ws_bytes);
  cudaStream_t st;
  cudaStreamCreate(&st);
  int rounds = 0;
  const int rc = b200_rnnt_greedy(enc_proj, lens, B, T, embed, lstm, w_c, b_c, w_out, b_out, V, Emb, H, L, J,
                                  blank, n_steps, tokens, n_tok, max_tokens, ws, ws_bytes, &rounds,
                                  (b200_stream_t)st);
  if (rc != B200_OK) {
    fprintf(stderr, "b200_rnnt_greedy failed (%d): %s\n", rc, b200_last_error());
    return 5;
  }
  cudaStreamSynchronize(st);
  int32_t* h_n = (int32_t*)malloc(sizeof(int32_t) * (size_t)B);
  int32_t* h_t = (int32_t*)malloc(sizeof(int32_t) * (size_t)B * max_tokens);
  cudaMemcpy(h_n, n_tok, sizeof(int32_t) * (size_t)B, cudaMemcpyDeviceToHost);
  cudaMemcpy(h_t, tokens, sizeof(int32_t) * (size_t)B * max_tokens, cudaMemcpyDeviceToHost);
  FILE* o = fopen(argv[2], "wb");
  if (!o) return 6;
  int32_t r32 = rounds;
  fwrite(&r32, 4, 1, o);
  fwrite(h_n, 4, (size_t)B, o);
  fwrite(h_t, 4, (size_t)B * max_tokens, o);
  fclose(o);
  printf("decoded %d utterances in %d rounds, %llu kernel launches\n", B, rounds,
         (unsigned long long)b200_launch_count());
  return 0;
}
