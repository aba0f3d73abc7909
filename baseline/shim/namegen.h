// namegen.h — stand-in for the header the reference includes but does not ship
// (/root/reference/namegensf.cu:1).  It supplies exactly what that file uses:
// six model macros, the 27 OFFSETk float offsets (ref:375-407) and the three
// prototypes (ref:359, :627, :897).  HIDDEN_DIM and NUM_CHAR are confirmed by
// comments in the reference (ref:694, :862); the other four are the course
// skeleton's values and can be overridden with -D.
#pragma once

#ifndef NUM_CHAR
#define NUM_CHAR 256
#endif
#ifndef EMBEDDING_DIM
#define EMBEDDING_DIM 512
#endif
#ifndef HIDDEN_DIM
#define HIDDEN_DIM 1024
#endif
#ifndef MAX_LEN
#define MAX_LEN 10
#endif
#ifndef SOS
#define SOS 0
#endif
#ifndef EOS
#define EOS 1
#endif

// contiguous layout in OFFSET order
#define SZ_EMB (NUM_CHAR * EMBEDDING_DIM)
#define SZ_WI0 (HIDDEN_DIM * EMBEDDING_DIM)
#define SZ_WH (HIDDEN_DIM * HIDDEN_DIM)
#define SZ_B (HIDDEN_DIM)

#define OFFSET0 0
#define OFFSET1 (OFFSET0 + SZ_EMB)
#define OFFSET2 (OFFSET1 + SZ_WI0)
#define OFFSET3 (OFFSET2 + SZ_WI0)
#define OFFSET4 (OFFSET3 + SZ_WI0)
#define OFFSET5 (OFFSET4 + SZ_WH)
#define OFFSET6 (OFFSET5 + SZ_WH)
#define OFFSET7 (OFFSET6 + SZ_WH)
#define OFFSET8 (OFFSET7 + SZ_WH)
#define OFFSET9 (OFFSET8 + SZ_WH)
#define OFFSET10 (OFFSET9 + SZ_WH)
#define OFFSET11 (OFFSET10 + SZ_WH)
#define OFFSET12 (OFFSET11 + SZ_WH)
#define OFFSET13 (OFFSET12 + SZ_WH)
#define OFFSET14 (OFFSET13 + SZ_B)
#define OFFSET15 (OFFSET14 + SZ_B)
#define OFFSET16 (OFFSET15 + SZ_B)
#define OFFSET17 (OFFSET16 + SZ_B)
#define OFFSET18 (OFFSET17 + SZ_B)
#define OFFSET19 (OFFSET18 + SZ_B)
#define OFFSET20 (OFFSET19 + SZ_B)
#define OFFSET21 (OFFSET20 + SZ_B)
#define OFFSET22 (OFFSET21 + SZ_B)
#define OFFSET23 (OFFSET22 + SZ_B)
#define OFFSET24 (OFFSET23 + SZ_B)
#define OFFSET25 (OFFSET24 + SZ_B)
#define OFFSET26 (OFFSET25 + NUM_CHAR * HIDDEN_DIM)
#define PARAM_FLOATS (OFFSET26 + NUM_CHAR)

void namegen_initialize(int N, int rng_seed, char *parameter_fname);
void namegen(int N, float *random_floats, char *output);
void namegen_finalize();
