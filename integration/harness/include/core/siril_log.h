/* TEST HARNESS ONLY — src/core/siril_log.h:30 */
#ifndef HARNESS_LOG_H
#define HARNESS_LOG_H
char *siril_log_message(const char *format, ...);
#define siril_log_error siril_log_message
#define siril_log_debug(...) do { } while (0)
#endif
