/* Key columns of TPC-H customer / orders / lineitem from the reference's own dbgen (third_party/tpch-dbgen), compiled from
 * the sources where they lie into oracle/_ref/libref_dbgen.so. Test infrastructure: produces the C1 / C4 join inputs
 * "as dbgen makes them" (SURVEY.md 8d). The call pattern restates TpchDbGenerator::generate()
 * (src/benchmarklib/tpch/tpch_db_generator.cpp:203-237): dbgen_reset_seeds(); customers 1..n with
 * row_start(CUST); mk_cust(); row_stop(CUST); then orders 1..n with row_start(ORDER); mk_order(i, &o, 0, scale);
 * row_stop(ORDER); every order carries its lineitems. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <dss.h>
#include <dsstypes.h>
#include <rnd.h>

void dbgen_reset_seeds();

/* row counts at a scale factor: customer, orders (tdefs[].base * scale, tpch_db_generator.cpp:208,219) */
void ref_dbgen_counts(float scale, int64_t* n_customer, int64_t* n_orders) {
  *n_customer = (int64_t)(tdefs[CUST].base * scale);
  *n_orders = (int64_t)(tdefs[ORDER].base * scale);
}

/* Fills c_custkey[n_customer], o_orderkey[n_orders], o_custkey[n_orders], o_lines[n_orders] and, when l_orderkey is not
 * NULL, l_orderkey[*n_lineitem] (capacity l_cap). Returns 0, or -1 if l_cap is too small (n_lineitem holds the need). */
int ref_dbgen_keys(float scale, int32_t* c_custkey, int32_t* o_orderkey, int32_t* o_custkey, int32_t* o_lines,
                   int32_t* l_orderkey, int64_t l_cap, int64_t* n_lineitem) {
  int64_t n_cust, n_ord, n_line = 0;
  int rc = 0;
  ref_dbgen_counts(scale, &n_cust, &n_ord);
  dbgen_reset_seeds();
  for (int64_t i = 0; i < n_cust; ++i) {
    customer_t c;
    memset(&c, 0, sizeof(c));
    row_start(CUST);
    mk_cust(i + 1, &c);
    row_stop(CUST);
    if (c_custkey) c_custkey[i] = (int32_t)c.custkey;
  }
  for (int64_t i = 0; i < n_ord; ++i) {
    order_t o;
    memset(&o, 0, sizeof(o));
    row_start(ORDER);
    mk_order(i + 1, &o, 0, scale);
    row_stop(ORDER);
    if (o_orderkey) o_orderkey[i] = (int32_t)o.okey;
    if (o_custkey) o_custkey[i] = (int32_t)o.custkey;
    if (o_lines) o_lines[i] = (int32_t)o.lines;
    for (long l = 0; l < o.lines; ++l) {
      if (l_orderkey) {
        if (n_line < l_cap) l_orderkey[n_line] = (int32_t)o.l[l].okey; else rc = -1;
      }
      ++n_line;
    }
  }
  *n_lineitem = n_line;
  return rc;
}

/* o_custkey alone, for the large scale factors the full generator is too slow for (C4 at SF100: 150 M orders with their
 * comments and lineitems take minutes). mk_order draws the customer key with ONE value of random stream O_CKEY_SD per order
 * (build.c:180-189) and row_stop(ORDER) moves every stream to its per-row boundary, which is 1 for that stream (the Seed[]
 * table in rnd.h / driver.c): the column IS that stream. The draw and the CUST_MORTALITY loop below are build.c's own
 * statements, run on the reference's RANDOM; tests/test_dbgen_keys.py checks the result against mk_order's at SF 0.1 and 1. */
int ref_dbgen_o_custkey(float scale, int32_t* o_custkey) {
  int64_t n_cust, n_ord;
  ref_dbgen_counts(scale, &n_cust, &n_ord);
  if (scale >= 30000) return -1; /* RANDOM64 territory: not needed here */
  dbgen_reset_seeds();
  for (int64_t i = 0; i < n_ord; ++i) {
    DSS_HUGE custkey;
    int delta = 1;
    row_start(ORDER);
    RANDOM(custkey, O_CKEY_MIN, O_CKEY_MAX, O_CKEY_SD);
    while (custkey % CUST_MORTALITY == 0) {
      custkey += delta;
      custkey = MIN(custkey, O_CKEY_MAX);
      delta *= -1;
    }
    /* row_stop(ORDER) for this stream alone: usage 1 == boundary 1, nothing to skip */
    o_custkey[i] = (int32_t)custkey;
  }
  return 0;
}
