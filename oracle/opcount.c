/* oracle/opcount.c -- exact count of the FP64 operations the REFERENCE executes per point
 * (TEST / MEASUREMENT INFRASTRUCTURE, not part of the product).
 *
 * SURVEY.md section 8(d) asks for algorithmic flop counts "from an op-counting build of the oracle
 * math".  The reference's sources use plain `double` (no scalar-type template), so instead of
 * re-typing them this tool counts what the UNMODIFIED reference binary (oracle/_ref/
 * libusgscsm_ref.so) actually executes: the per-point loop runs in a forked child under
 * ptrace(PTRACE_SINGLESTEP); the parent histograms every instruction address that falls inside
 * the library, and counts calls of the libm entry points (sin, cos, ...).  oracle/opcount.py maps
 * the addresses to mnemonics with objdump and applies the counting convention of SURVEY.md 8(d)
 * (add / sub / mul / div / sqrt / compare / abs = 1 each, sin or cos = 1, integer work = 0).
 *
 *   opcount <lib.so> <state file> <g2i|i2g> <points.bin> <n> <prec> <out.txt>
 * points.bin: n x 3 doubles, SoA (x[n] y[n] z[n]  or  line[n] samp[n] height[n]).
 */
#define _GNU_SOURCE
#include <dlfcn.h>
#include <link.h>
#include <signal.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/ptrace.h>
#include <sys/user.h>
#include <sys/wait.h>
#include <unistd.h>

typedef void *(*from_state_fn)(const char *, size_t);
typedef int (*g2i_fn)(void *, long, const double *, const double *, const double *, double,
                      double *, double *, double *, uint32_t *, int);
typedef int (*i2g_fn)(void *, long, const double *, const double *, const double *, int, double,
                      double *, double *, double *, double *, uint32_t *, int);

static const char *g_libname;
static uintptr_t g_base, g_lo = ~(uintptr_t)0, g_hi;

static int phdr_cb(struct dl_phdr_info *info, size_t size, void *data) {
  (void)size;
  (void)data;
  if (!info->dlpi_name || !strstr(info->dlpi_name, g_libname)) return 0;
  g_base = info->dlpi_addr;
  for (int i = 0; i < info->dlpi_phnum; i++) {
    const ElfW(Phdr) *p = &info->dlpi_phdr[i];
    if (p->p_type != PT_LOAD || !(p->p_flags & PF_X)) continue;
    uintptr_t lo = info->dlpi_addr + p->p_vaddr, hi = lo + p->p_memsz;
    if (lo < g_lo) g_lo = lo;
    if (hi > g_hi) g_hi = hi;
  }
  return 1;
}

int main(int argc, char **argv) {
  if (argc != 8) {
    fprintf(stderr, "usage: opcount lib.so state g2i|i2g points.bin n prec out.txt\n");
    return 2;
  }
  const char *lib = argv[1], *dir = argv[3];
  long n = atol(argv[5]);
  double prec = atof(argv[6]);
  void *h = dlopen(lib, RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    fprintf(stderr, "dlopen: %s\n", dlerror());
    return 1;
  }
  const char *slash = strrchr(lib, '/');
  g_libname = slash ? slash + 1 : lib;
  dl_iterate_phdr(phdr_cb, NULL);
  if (!g_hi) {
    fprintf(stderr, "library text segment not found\n");
    return 1;
  }
  FILE *f = fopen(argv[2], "rb");
  if (!f) return 1;
  fseek(f, 0, SEEK_END);
  long len = ftell(f);
  fseek(f, 0, SEEK_SET);
  char *state = malloc(len + 1);
  if (fread(state, 1, len, f) != (size_t)len) return 1;
  fclose(f);
  from_state_fn from_state = (from_state_fn)dlsym(h, "ref_model_from_state");
  void *model = from_state(state, (size_t)len);
  if (!model) {
    fprintf(stderr, "model: %s\n", ((const char *(*)(void))dlsym(h, "ref_last_error"))());
    return 1;
  }
  double *in = malloc(sizeof(double) * 3 * n), *out = malloc(sizeof(double) * 4 * n);
  uint32_t *st = malloc(sizeof(uint32_t) * n);
  f = fopen(argv[4], "rb");
  if (!f || fread(in, sizeof(double), 3 * n, f) != (size_t)(3 * n)) return 1;
  fclose(f);
  g2i_fn g2i = (g2i_fn)dlsym(h, "ref_ground_to_image");
  i2g_fn i2g = (i2g_fn)dlsym(h, "ref_image_to_ground");
  /* libm entry points whose CALLS are counted (their internals are not FP64 "flops" of the model) */
  const char *mnames[] = {"sin", "cos", "sincos", "tan", "atan", "atan2", "acos", "asin", "pow",
                          "exp", "log", "fmod", "floor", "ceil", "round", "sqrt"};
  enum { NM = sizeof(mnames) / sizeof(mnames[0]) };
  uintptr_t maddr[NM];
  unsigned long mcount[NM];
  for (int i = 0; i < NM; i++) {
    maddr[i] = (uintptr_t)dlsym(RTLD_DEFAULT, mnames[i]);
    mcount[i] = 0;
  }
  /* warm the lazy PLT bindings and the allocator with one untraced pass over a few points */
  long w = n < 4 ? n : 4;
  if (!strcmp(dir, "g2i")) g2i(model, w, in, in + n, in + 2 * n, prec, out, out + n, out + 2 * n, st, 1);
  else i2g(model, w, in, in + n, in + 2 * n, 1, prec, out, out + n, out + 2 * n, out + 3 * n, st, 1);

  pid_t pid = fork();
  if (pid == 0) {
    ptrace(PTRACE_TRACEME, 0, 0, 0);
    raise(SIGSTOP);
    if (!strcmp(dir, "g2i")) g2i(model, n, in, in + n, in + 2 * n, prec, out, out + n, out + 2 * n, st, 1);
    else i2g(model, n, in, in + n, in + 2 * n, 1, prec, out, out + n, out + 2 * n, out + 3 * n, st, 1);
    _exit(0);
  }
  int status;
  waitpid(pid, &status, 0);
  size_t span = g_hi - g_lo;
  uint32_t *hist = calloc(span, sizeof(uint32_t));
  unsigned long steps = 0, inside = 0;
  struct user_regs_struct regs;
  for (;;) {
    if (ptrace(PTRACE_SINGLESTEP, pid, 0, 0) < 0) break;
    waitpid(pid, &status, 0);
    if (WIFEXITED(status) || WIFSIGNALED(status)) break;
    if (ptrace(PTRACE_GETREGS, pid, 0, &regs) < 0) break;
    uintptr_t ip = regs.rip;
    steps++;
    if (ip >= g_lo && ip < g_hi) {
      hist[ip - g_lo]++;
      inside++;
    } else {
      for (int i = 0; i < NM; i++)
        if (ip == maddr[i]) mcount[i]++;
    }
  }
  f = fopen(argv[7], "w");
  fprintf(f, "# points %ld direction %s prec %g steps %lu inside_library %lu\n", n, dir, prec, steps, inside);
  for (int i = 0; i < NM; i++)
    if (mcount[i]) fprintf(f, "libm %s %lu\n", mnames[i], mcount[i]);
  for (size_t i = 0; i < span; i++)
    if (hist[i]) fprintf(f, "%lx %u\n", (unsigned long)(g_lo + i - g_base), hist[i]);
  fclose(f);
  return 0;
}
