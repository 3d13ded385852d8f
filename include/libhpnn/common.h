/*
 * libhpnn/common.h -- compatibility substrate of hpnn-b200.
 *
 * Host programs written against libhpnn (the reference's own tests/train_nn.c,
 * tests/run_nn.c and tutorial converters among them) use a vocabulary of type aliases
 * and helper macros from this header.  This file provides that vocabulary -- same names,
 * same observable behaviour (reference: include/libhpnn/common.h:60-330,580-605) -- on
 * top of libc only.  The CUDA_* / CUBLAS_* macro families of the reference are NOT
 * reproduced: device work goes through the hpnn_core C ABI, never through macros.
 */
#ifndef COMMON_H
#define COMMON_H

#include <ctype.h>
#include <dirent.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/types.h>
#include <unistd.h>

#include "unroll.def"

/* ---- scalar vocabulary --------------------------------------------------- */
#define DIR_S DIR
#define CHAR char
#define UCHAR unsigned char
#define SHORT short
#define UINT unsigned int
#define UINT64 uint64_t
#define DOUBLE double
#define BOOL int
#ifndef TRUE
#define TRUE (1 == 1)
#endif
#ifndef FALSE
#define FALSE (1 == 0)
#endif
#define TINY 1E-14

#if defined(__GNUC__)
#define FUNCTION __PRETTY_FUNCTION__
#else
#define FUNCTION __func__
#endif
#define QUOTE(a) #a

/* ---- output: one writer, rank 0 only when launched by torchrun/mpirun -------- */
#ifdef __cplusplus
extern "C" {
#endif
int hpnn_compat_is_root(void); /* 1 unless RANK / OMPI_COMM_WORLD_RANK / PMI_RANK says otherwise */
#ifdef __cplusplus
}
#endif
#define _OUT(_file, ...)                                        \
  do {                                                          \
    if (hpnn_compat_is_root()) fprintf((_file), __VA_ARGS__);   \
  } while (0)

/* ---- character / number helpers -------------------------------------------- */
#define STRFIND(a, b) strstr(b, a)
#define ISDIGIT(a) isdigit((unsigned char)(a))
#define ISGRAPH(a) isgraph((unsigned char)(a))
#define ISSPACE(a) isspace((unsigned char)(a))
#define STR2ULL strtoull
#define STR2D strtod

/* ---- memory: zero-filled or die -------------------------------------------- */
#define ALLOC(pointer, size, type)                                                      \
  do {                                                                                  \
    pointer = (type *)calloc((size), sizeof(type));                                     \
    if (pointer == NULL) {                                                              \
      _OUT(stderr, "Alloc error (function %s, line %i)\n", FUNCTION, __LINE__);         \
      exit(-1);                                                                         \
    }                                                                                   \
  } while (0)
#define ALLOC_REPORT(pointer, size, type, mem) \
  do {                                         \
    ALLOC(pointer, size, type);                \
    mem += (size) * sizeof(type);              \
  } while (0)
#define FREE(pointer)                      \
  do {                                     \
    if (pointer != NULL) free(pointer);    \
    pointer = NULL;                        \
  } while (0)

/* ---- strings ------------------------------------------------------------------ */
#define STRLEN(src, len) \
  while ((len < UINT_LEAST32_MAX) && (src[len] != '\0')) len++
/* dest = fresh copy of src; an empty src gives dest = NULL (reference behaviour) */
#define STRDUP(src, dest)                     \
  do {                                        \
    size_t hp_n_ = strlen(src);               \
    if (hp_n_ != 0) {                         \
      ALLOC(dest, hp_n_ + 1, CHAR);           \
      memcpy(dest, src, hp_n_);               \
    } else {                                  \
      dest = NULL;                            \
    }                                         \
  } while (0)
#define STRDUP_REPORT(src, dest, mem)         \
  do {                                        \
    size_t hp_n_ = strlen(src);               \
    if (hp_n_ != 0) {                         \
      ALLOC(dest, hp_n_ + 1, CHAR);           \
      memcpy(dest, src, hp_n_);               \
      mem += hp_n_ * sizeof(CHAR);            \
    } else {                                  \
      dest = NULL;                            \
    }                                         \
  } while (0)
/* dest = src1 ++ src2; NULL when src2 is empty (reference behaviour) */
#define STRCAT(dest, src1, src2)                       \
  do {                                                 \
    size_t hp_a_ = strlen(src1), hp_b_ = strlen(src2); \
    if ((hp_a_ + hp_b_ > 0) && (hp_b_ > 0)) {          \
      ALLOC(dest, hp_a_ + hp_b_ + 1, CHAR);            \
      memcpy(dest, src1, hp_a_);                       \
      memcpy(dest + hp_a_, src2, hp_b_);               \
    } else {                                           \
      dest = NULL;                                     \
    }                                                  \
  } while (0)

/* ---- line-oriented parsing ------------------------------------------------------ */
#define PREP_READLINE() size_t _readline_len = 0
#define READLINE(fp, buffer)                                   \
  do {                                                         \
    ssize_t hp_rc_ = getline(&buffer, &_readline_len, fp);     \
    (void)hp_rc_;                                              \
  } while (0)
#define SKIP_BLANK(pointer) \
  while ((!ISGRAPH(*pointer)) && (*pointer != '\n') && (*pointer != '\0')) pointer++
#define SKIP_NUM(pointer) \
  while ((ISDIGIT(*pointer)) && (*pointer != '\n') && (*pointer != '\0')) pointer++
/* cut the string at the first blank, tab, newline or '#' */
#define STR_CLEAN(pointer)                                                         \
  do {                                                                             \
    CHAR *hp_p_ = pointer;                                                         \
    while (*hp_p_ != '\0') {                                                       \
      if (*hp_p_ == '\t' || *hp_p_ == ' ' || *hp_p_ == '\n' || *hp_p_ == '#') {    \
        *hp_p_ = '\0';                                                             \
        break;                                                                     \
      }                                                                            \
      hp_p_++;                                                                     \
    }                                                                              \
  } while (0)
#define GET_LAST_LINE(fp, buffer)                       \
  do {                                                  \
    fseek(fp, -2, SEEK_END);                            \
    while (fgetc(fp) != '\n') fseek(fp, -2, SEEK_CUR);  \
    fseek(fp, +1, SEEK_CUR);                            \
    READLINE(fp, buffer);                               \
  } while (0)
#define GET_UINT(i, in, out)              \
  do {                                    \
    i = (UINT)STR2ULL(in, &(out), 10);    \
  } while (0)
#define GET_DOUBLE(d, in, out)      \
  do {                              \
    d = (DOUBLE)STR2D(in, &(out));  \
  } while (0)
#define ARRAY_CP(src, dest, n)                                  \
  do {                                                          \
    if (src != NULL) {                                          \
      UINT hp_i_;                                               \
      for (hp_i_ = 0; hp_i_ < (n); hp_i_++) dest[hp_i_] = src[hp_i_]; \
    }                                                           \
  } while (0)
#define ASSERTPTR(pointer, retval)                                                   \
  do {                                                                               \
    if (pointer == NULL) {                                                           \
      _OUT(stderr, "Error: NULL pointer (function %s, line %i):\n%s=NULL\n", FUNCTION, \
           __LINE__, QUOTE(pointer));                                                \
      return retval;                                                                 \
    }                                                                                \
  } while (0)
#define ASSERT_GOTO(pointer, label)                                                  \
  do {                                                                               \
    if (pointer == NULL) {                                                           \
      _OUT(stderr, "Error: NULL pointer (function %s, line %i):\n%s=NULL\n", FUNCTION, \
           __LINE__, QUOTE(pointer));                                                \
      goto label;                                                                    \
    }                                                                                \
  } while (0)

/* ---- directories ------------------------------------------------------------------- */
#define GET_CWD(cwd)        \
  do {                      \
    cwd = getcwd(NULL, 0);  \
  } while (0)
#define OPEN_DIR(dir, path) \
  do {                      \
    dir = opendir(path);    \
  } while (0)
#define FILE_FROM_DIR(dir, file)             \
  do {                                       \
    struct dirent *hp_e_ = readdir(dir);     \
    if (hp_e_ == NULL) file = NULL;          \
    else STRDUP(hp_e_->d_name, file);        \
  } while (0)
#define CLOSE_DIR(dir, ok) \
  do {                     \
    ok = closedir(dir);    \
  } while (0)

/* ---- multi-GPU descriptors kept for ABI compatibility -------------------------------- */
/* The reference elects one of three memory models (src/libhpnn.c:245-280).  hpnn-b200 has a
 * single model -- weights partitioned across GPUs, small symmetric buffers mapped by every
 * peer -- and reports it as CUDA_MEM_P2P (what the reference itself elects on NVSwitch). */
typedef enum {
  CUDA_MEM_NONE, /* one GPU */
  CUDA_MEM_EXP,  /* reference: full replica per GPU + explicit copies (never used here) */
  CUDA_MEM_P2P,  /* peer access; here: partitioned weights + peer stores */
  CUDA_MEM_CMM,  /* reference: managed memory (never used here) */
} cudas_mem;

typedef struct {
  UINT n_gpu;
  void *cuda_handle;   /* reference: cublasHandle_t[] -- always NULL, no library GEMV here */
  UINT cuda_n_streams; /* accepted and stored; the persistent kernel needs one stream */
  void *cuda_streams;  /* always NULL */
  cudas_mem mem_model;
} cudastreams;

#endif /* COMMON_H */
