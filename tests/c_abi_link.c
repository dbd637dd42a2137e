/* Compile-and-link check: include/oe_abi.h is plain C, and a C caller can name every entry point.  Built (not run) by
 * tests/test_abi_exports.py with `gcc -std=c99 -Wall -Werror`. */
#include <stddef.h>
#include "../include/oe_abi.h"

int main(void) {
  oe_config cfg = {0, 0, 1, OE_CFG_DEFAULT, NULL};
  oe_engine* h = NULL;
  oe_stats st;
  int64_t n = 0;
  if (oe_abi_version() != OE_ABI_VERSION) return 2;
  if (oe_create(&cfg, &h) != OE_OK) return oe_last_error(NULL) == NULL;   /* no GPU on the build box: a clean refusal */
  /* never reached without a device; the calls below only have to link */
  (void)oe_load_programs(h, NULL); (void)oe_load_tasks(h, NULL); (void)oe_load_offers(h, NULL); (void)oe_load_requirements(h, NULL);
  (void)oe_matrix_stride_words(h); (void)oe_feasibility(h, NULL, &st); (void)oe_class_feasibility(h, NULL, &st);
  (void)oe_evaluate_cycle(h, NULL, &st); (void)oe_stage_failures(h, 0, 0, NULL); (void)oe_unexpected_resources(h, NULL, 0, NULL, &n);
  (void)oe_update_offers(h, NULL, 0, NULL); (void)oe_apply_launches(h); (void)oe_export_tasks(h, &n, NULL, NULL, NULL, NULL, NULL, NULL, NULL, NULL, NULL, NULL);
  (void)oe_device_buffer(h, OE_BUF_DECISIONS, NULL, &n); (void)oe_stream(h);
  (void)oe_first_fit_static(0, NULL, NULL, 0, 0, 0, NULL, 0, NULL, NULL, 0);
  oe_destroy(h);
  return 0;
}
