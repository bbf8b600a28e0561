#include "../../include/rfq_parse.h"
#include "ctx.h"
extern "C" int rfq_csv_run(rfq_ctx* c, const int64_t*) {
  if (c) c->err = "csv path not built yet";
  return RFQ_ERR_STATE;
}
