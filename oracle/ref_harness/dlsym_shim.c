/* dlsym_shim.c — glibc >= 2.34 no longer exports _dl_sym@GLIBC_PRIVATE, which the reference
 * libvgpu.so (built against an older glibc) calls in preInit (libvgpu.so@0x16c94).  An
 * unversioned definition satisfies it.  TEST INFRASTRUCTURE ONLY. */
#define _GNU_SOURCE
#include <dlfcn.h>
void *_dl_sym(void *handle, const char *name, void *who) {
    (void)handle; (void)who;
    void *libc = dlopen("libc.so.6", RTLD_LAZY | RTLD_NOLOAD);
    return dlvsym(libc, name, "GLIBC_2.2.5");
}
