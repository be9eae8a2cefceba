/* Test infrastructure: a minimal NVTX v3 injection library (what a profiler installs through
 * NVTX_INJECTION64_PATH).  It hooks nvtxRangePushA / nvtxRangePop of the core module and appends one line per
 * call to the file named by FEMTC_NVTX_LOG, so that tests/test_abi.py can show that libfemtc_b200.so opens and
 * closes its NVTX ranges (the stand-in for the reference's LIKWID marker regions, bench_mass.cpp:15-24,94,108-109)
 * without nsys on the box.  Protocol: nvtx3/nvtxDetail/nvtxInit.h, nvtxTypes.h (ids restated below). */
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>

typedef void (*fnptr)(void);
typedef fnptr** function_table;
typedef struct
{
  size_t struct_size;
  int (*GetModuleFunctionTable)(int module, function_table* out_table, unsigned int* out_size);
} export_table_callbacks;
typedef const void* (*get_export_table_fn)(unsigned int export_table_id);

enum { ETID_CALLBACKS = 1, CB_MODULE_CORE = 1, CBID_CORE_RangePushA = 9, CBID_CORE_RangePop = 11 };

static void log_line(const char* kind, const char* msg)
{
  const char* path = getenv("FEMTC_NVTX_LOG");
  if (!path)
    return;
  FILE* f = fopen(path, "a");
  if (!f)
    return;
  fprintf(f, "%s %s\n", kind, msg ? msg : "");
  fclose(f);
}
static int depth = 0;
static int my_push(const char* message)
{
  log_line("push", message);
  return depth++;
}
static int my_pop(void)
{
  log_line("pop", "");
  return --depth;
}

int InitializeInjectionNvtx2(get_export_table_fn get_export_table)
{
  const export_table_callbacks* cb = (const export_table_callbacks*)get_export_table(ETID_CALLBACKS);
  function_table table = NULL;
  unsigned int size = 0;
  if (!cb || !cb->GetModuleFunctionTable(CB_MODULE_CORE, &table, &size) || size <= CBID_CORE_RangePop)
    return 0;
  *table[CBID_CORE_RangePushA] = (fnptr)my_push;
  *table[CBID_CORE_RangePop] = (fnptr)my_pop;
  log_line("init", "femtc nvtx probe");
  return 1;
}
