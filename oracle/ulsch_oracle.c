/* TEST INFRASTRUCTURE - not part of the product. Plain-C restatement of the reference's UL-SCH demultiplexer
 * (TS 38.212 6.2.7): lib/phy/upper/channel_processors/ulsch_demultiplex_impl.cpp of srsRAN Project 23.5. It walks the
 * resource elements one by one exactly as the reference does (the product computes the same mapping in closed form per
 * element, srsran_project_23.5_b200/csrc/ldpc_ulsch.cuh) and is pinned against the reference's own class through
 * oracle/ref_shim.cpp (tests/test_oracle_vs_reference.py). Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may
 * use it. */
#include <stdint.h>
#include <string.h>

typedef struct {
  unsigned mod_order, nof_layers, nof_prb, start_symbol, nof_symbols, dmrs_type, nof_cdm_groups_without_data;
  unsigned dmrs_symbol_mask, nof_harq_ack_rvd;
} oracle_ulsch_config;

/* What happens to one resource element; out-of-band consumer of the walk. */
typedef struct {
  /* demultiplex mode */
  const int8_t* in;
  int8_t *      sch, *ack, *csi1, *csi2;
  unsigned      bpr;
  /* placeholder mode */
  int       placeholders;
  unsigned  one_ack, one_csi1, one_csi2; /* the field carries exactly one information bit */
  unsigned  re_counter;
  uint16_t* list;
  unsigned  list_len, list_max;
} walk_sink;

static void take(walk_sink* s, int8_t** dst, int zero, unsigned one_bit)
{
  if (s->placeholders) {
    /* get_placeholders lambdas, ulsch_demultiplex_impl.cpp:411-445: zero-filled elements neither count nor list */
    if (zero) {
      return;
    }
    if (one_bit) {
      if (s->list_len < s->list_max) {
        s->list[s->list_len] = (uint16_t)s->re_counter;
      }
      s->list_len++;
    }
    s->re_counter++;
    return;
  }
  /* demultiplex lambdas, :335-374 */
  if (zero) {
    memset(*dst, 0, s->bpr);
  } else {
    memcpy(*dst, s->in, s->bpr);
    s->in += s->bpr;
  }
  *dst += s->bpr;
}

static unsigned ceil_div(unsigned a, unsigned b)
{
  return (a + b - 1) / b;
}

/* ulsch_demultiplex_generic, ulsch_demultiplex_impl.cpp:76-312. Returns 0, or -1 where the reference asserts. */
static int walk(const oracle_ulsch_config* c, unsigned G_ack, unsigned G_csi1, unsigned G_csi2, walk_sink* sink)
{
  const unsigned bpr = c->mod_order * c->nof_layers;
  int            first_dmrs = -1, l1 = -1, l1_csi = -1, l;
  for (l = 0; l != 14; ++l) { /* :33-55 */
    const int d = (c->dmrs_symbol_mask >> l) & 1;
    if (d && first_dmrs < 0) {
      first_dmrs = l;
    }
    if (!d && first_dmrs >= 0 && l1 < 0) {
      l1 = l;
    }
    if (!d && l1_csi < 0) {
      l1_csi = l;
    }
  }
  if (first_dmrs < 0 || l1 < 0 || l1_csi < 0 || bpr == 0) {
    return -1;
  }
  /* :57-74 */
  const unsigned re_per_prb_dmrs = c->nof_cdm_groups_without_data * (c->dmrs_type == 1 ? 6u : 4u);
  if (re_per_prb_dmrs > 12) {
    return -1;
  }
  const unsigned nof_re_dmrs = (12 - re_per_prb_dmrs) * c->nof_prb;
  const unsigned G_rvd       = c->nof_harq_ack_rvd;
  unsigned       m_rvd = 0, m_ack = 0, m_csi1 = 0, m_csi2 = 0;
  unsigned       sym;
  for (sym = c->start_symbol; sym != c->start_symbol + c->nof_symbols; ++sym) {
    unsigned i;
    if ((c->dmrs_symbol_mask >> sym) & 1) { /* :126-131 */
      for (i = 0; i != nof_re_dmrs; ++i) {
        take(sink, &sink->sch, 0, 0);
      }
      continue;
    }
    const unsigned M_sc  = c->nof_prb * 12;
    unsigned       M_uci = M_sc, M_rvd = 0;
    unsigned       ack_d = 0, ack_n = 0, rvd_d = 0, rvd_n = 0, csi1_d = 0, csi1_n = 0, csi2_d = 0, csi2_n = 0;
    if ((int)sym >= l1) { /* :154-188 */
      const unsigned rvd_left = G_rvd - m_rvd, ack_left = G_ack - m_ack;
      if (G_rvd != 0 && rvd_left != 0) {
        rvd_d = 1, rvd_n = M_uci;
        if (rvd_left < M_uci * bpr) {
          rvd_d = (M_uci * bpr) / rvd_left;
          rvd_n = ceil_div(rvd_left, bpr);
        }
        M_rvd = rvd_n;
        if (ack_left != 0) {
          ack_d = 1, ack_n = M_rvd;
          if (ack_left < M_rvd * bpr) {
            ack_d = (M_rvd * bpr) / ack_left;
            ack_n = ceil_div(ack_left, bpr);
          }
        }
      } else if (ack_left != 0) {
        ack_d = 1, ack_n = M_uci;
        if (ack_left < M_uci * bpr) {
          ack_d = (M_uci * bpr) / ack_left;
          ack_n = ceil_div(ack_left, bpr);
        }
        M_uci -= ack_n;
      }
    }
    if ((int)sym >= l1_csi) { /* :191-219 */
      const unsigned csi1_left = G_csi1 - m_csi1, csi2_left = G_csi2 - m_csi2;
      if (M_uci > M_rvd && csi1_left != 0) {
        csi1_d = 1, csi1_n = M_uci - M_rvd;
        if (csi1_left < (M_uci - M_rvd) * bpr) {
          csi1_d = ((M_uci - M_rvd) * bpr) / csi1_left;
          csi1_n = ceil_div(csi1_left, bpr);
        }
        M_uci -= csi1_n;
      }
      if (M_uci > 0 && csi2_left != 0) {
        csi2_d = 1, csi2_n = M_uci;
        if (csi2_left < M_uci * bpr) {
          csi2_d = (M_uci * bpr) / csi2_left;
          csi2_n = ceil_div(csi2_left, bpr);
        }
        M_uci -= csi2_n;
      }
    }
    unsigned sch_n = M_uci; /* :222 */
    m_rvd += rvd_n * bpr, m_ack += ack_n * bpr, m_csi1 += csi1_n * bpr, m_csi2 += csi2_n * bpr;
    unsigned i_ack = 0, i_csi1 = 0, i_csi2 = 0;
    for (i = 0; i != M_sc; ++i) { /* :231-300 */
      const int reserved = rvd_n != 0 && i % rvd_d == 0;
      int       zero     = 0;
      if (reserved) {
        --rvd_n;
      }
      if (G_rvd != 0) {
        if (reserved && ack_n != 0 && (i_ack++ % ack_d) == 0) {
          take(sink, &sink->ack, 0, sink->one_ack);
          --ack_n;
          zero = 1;
        }
      } else if (ack_n != 0 && (i_ack++ % ack_d) == 0) {
        take(sink, &sink->ack, 0, sink->one_ack);
        --ack_n;
        continue;
      }
      if (!reserved && csi1_n != 0 && (i_csi1++ % csi1_d) == 0) {
        take(sink, &sink->csi1, 0, sink->one_csi1);
        --csi1_n;
        continue;
      }
      if (csi2_n != 0 && (i_csi2++ % csi2_d) == 0) {
        take(sink, &sink->csi2, zero, sink->one_csi2);
        --csi2_n;
        continue;
      }
      if (sch_n == 0) {
        return -1;
      }
      take(sink, &sink->sch, zero, 0);
      --sch_n;
    }
    if (ack_n || csi1_n || csi2_n || sch_n) { /* :303-306 */
      return -1;
    }
  }
  return (m_rvd == G_rvd && m_ack == G_ack && m_csi1 == G_csi1 && m_csi2 == G_csi2) ? 0 : -1; /* :309-315 */
}

/* ulsch_demultiplex_impl::demultiplex, :316-392. Sizes of the fields in soft bits; the caller sizes sch generously. */
int oracle_ulsch_demultiplex(const oracle_ulsch_config* c, unsigned G_ack, unsigned G_csi1, unsigned G_csi2, const int8_t* in,
                             unsigned in_len, int8_t* sch, int8_t* ack, int8_t* csi1, int8_t* csi2, unsigned* in_used,
                             unsigned* sch_len)
{
  if (G_ack == 0 && G_csi1 == 0 && G_csi2 == 0) { /* :324-331: plain copy */
    memcpy(sch, in, in_len);
    *in_used = *sch_len = in_len;
    return 0;
  }
  walk_sink s;
  memset(&s, 0, sizeof(s));
  s.in = in, s.sch = sch, s.ack = ack, s.csi1 = csi1, s.csi2 = csi2, s.bpr = c->mod_order * c->nof_layers;
  const int rc = walk(c, G_ack, G_csi1, G_csi2, &s);
  *in_used     = (unsigned)(s.in - in);
  *sch_len     = (unsigned)(s.sch - sch);
  return rc;
}

/* ulsch_demultiplex_impl::get_placeholders, :394-453. Returns the number of placeholder elements or -1. */
int oracle_ulsch_placeholders(const oracle_ulsch_config* c, unsigned O_ack, unsigned G_ack, unsigned O_csi1, unsigned G_csi1,
                              unsigned O_csi2, unsigned G_csi2, uint16_t* list, unsigned list_max)
{
  if (c->mod_order < 2 || (O_ack != 1 && O_csi1 != 1 && O_csi2 != 1)) {
    return 0;
  }
  walk_sink s;
  memset(&s, 0, sizeof(s));
  s.placeholders = 1, s.one_ack = O_ack == 1, s.one_csi1 = O_csi1 == 1, s.one_csi2 = O_csi2 == 1;
  s.list = list, s.list_max = list_max;
  if (walk(c, G_ack, G_csi1, G_csi2, &s) != 0) {
    return -1;
  }
  return (int)s.list_len;
}
