// Argument block of the fused SelectedUnits sampling kernel (su_sample.cu); mirrors as_su_sample_args of the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#ifndef AS_SU_SAMPLE_ARGS_DEFINED
#define AS_SU_SAMPLE_ARGS_DEFINED
typedef struct as_su_sample_args {
  const float *key, *c0, *c1, *mt, *wf1t, *bf1, *wq2t, *bq2, *wih_t, *whh_t;
  const float *ln_i_w, *ln_i_b, *ln_h_w, *ln_h_b, *ln_c_w, *ln_c_b;
  const int64_t* entity_num;
  const uint8_t* su_mask;
  const float* noise;
  float* logits;
  int64_t* units;
  int64_t* num;
  float* extra;
  float* emb;
  int* steps_row;
  int* ended_rows;   /* two device ints, 0 at launch, or NULL: [0] rows that have raised their end flag, [1] the latest end step;
                        once [0] == R a CTA leaves its loop as soon as it has computed [1] steps */
  int R, N1, S, extra_units;
  float temperature;
} as_su_sample_args;
#endif

namespace as {

struct SuSampleDev {
  const float* key;            // [R, N1, 32]
  const float* c0;             // [R, 256]
  const float* c1;             // [256]
  const float* mt;             // [256 k][256 j]   (Wq1 Wf2)^T
  const float* wf1t;           // [32][256]        embed_fc1.weight^T
  const float* bf1;            // [256]
  const float* wq2t;           // [256][32]        query_fc2.weight^T
  const float* bq2;            // [32]
  const float* wih_t;          // [32][128]
  const float* whh_t;          // [32][128]
  const float* ln_i_w; const float* ln_i_b;   // [128]
  const float* ln_h_w; const float* ln_h_b;   // [128]
  const float* ln_c_w; const float* ln_c_b;   // [32]
  const long long* entity_num; // [R]
  const unsigned char* su_mask;// [R] 1 = the sampled action type takes selected units
  const float* noise;          // [S, R, N1]
  float* logits;               // [R, S, N1]
  long long* units;            // [R, S]
  long long* num;              // [R]
  float* extra;                // [R, N1]
  float* emb;                  // [R, 32]  mean of the picked keys after the last step
  int* steps_row;              // [R]      1 + step at which the row raised its end flag (0: started ended, S: never)
  int* ended_rows;             // [2] device ints, 0 at launch: ended rows, latest end step; or nullptr (always S steps)
  int R, N1, S, extra_units;
  float temperature;
};

int su_sample(const SuSampleDev& p, cudaStream_t st);

}  // namespace as
