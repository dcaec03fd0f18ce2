#include "rdyb_error.h"

#include <stdarg.h>
#include <stdio.h>

#include "rdyb200.h"

static _Thread_local char rdyb_errbuf[1024] = "";

int rdyb_fail(int code, const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(rdyb_errbuf, sizeof(rdyb_errbuf), fmt, ap);
  va_end(ap);
  return code;
}

const char *rdyb_last_error(void) { return rdyb_errbuf; }

int rdyb_version(int *major, int *minor) {
  if (major) *major = 0;
  if (minor) *minor = 1;
  return RDYB_SUCCESS;
}
