/* d4b_dump.h -- tiny named-array container used by the test infrastructure.
 *
 * TEST INFRASTRUCTURE: written by oracle/ref_driver.c (the real reference,
 * compiled into oracle/_ref/) and read by tests/ (numpy, tests/d4bio.py) and
 * by oracle/d4b_oracle.c.  Not part of the product.
 *
 * Layout: magic "D4BDUMP1", then records
 *   u32 name_len | name bytes | u32 dtype (0 f64, 1 i32, 2 i64) | u32 ndim |
 *   u64 shape[ndim] | raw little-endian data
 */
#ifndef D4B_DUMP_H
#define D4B_DUMP_H
#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <stdlib.h>

enum { D4B_F64 = 0, D4B_I32 = 1, D4B_I64 = 2 };

static inline FILE *
d4b_dump_open (const char *path)
{
  FILE *f = fopen (path, "wb");
  if (!f) { perror (path); exit (2); }
  fwrite ("D4BDUMP1", 1, 8, f);
  return f;
}

static inline void
d4b_dump_put (FILE * f, const char *name, int dtype, int ndim,
              const uint64_t * shape, const void *data)
{
  static const size_t esz[3] = { 8, 4, 8 };
  uint32_t nl = (uint32_t) strlen (name), dt = (uint32_t) dtype, nd = (uint32_t) ndim;
  uint64_t n = 1;
  for (int i = 0; i < ndim; i++) n *= shape[i];
  fwrite (&nl, 4, 1, f);
  fwrite (name, 1, nl, f);
  fwrite (&dt, 4, 1, f);
  fwrite (&nd, 4, 1, f);
  fwrite (shape, 8, (size_t) ndim, f);
  if (n) fwrite (data, esz[dtype], (size_t) n, f);
}

static inline void
d4b_dump_put1 (FILE * f, const char *name, int dtype, uint64_t n, const void *data)
{
  d4b_dump_put (f, name, dtype, 1, &n, data);
}

static inline void
d4b_dump_scalar_i64 (FILE * f, const char *name, int64_t v)
{
  d4b_dump_put1 (f, name, D4B_I64, 1, &v);
}

static inline void
d4b_dump_scalar_f64 (FILE * f, const char *name, double v)
{
  d4b_dump_put1 (f, name, D4B_F64, 1, &v);
}
#endif
