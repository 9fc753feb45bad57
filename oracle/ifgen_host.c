/* ifgen_host.c - CPU side of the synthetic IF generator (TEST INFRASTRUCTURE, not product).
 *
 * The reference ships no IF recordings (test/testdata_download_link.txt lists URLs only), so
 * the oracle, the CPU legs of bench.py and its --impl reference arm run on IF made here from
 * tests/testkit/ifgen.h (the model is documented there).  Ranging codes come from the oracle's
 * own generator (port/port_code.c, pinned to src/sdrcode.c), so this library depends on nothing
 * of the product: a process that times the reference's CPU path maps only oracle/ libraries.
 */
#include <string.h>
#include "../tests/testkit/ifgen.h"
#include "port/port.h"

int oracle_ifgen_host(const void *vcfg, uint64_t k0, uint64_t n, signed char *out)
{
    static __thread signed char codes[IFGEN_MAXSAT * IFGEN_MAXCLEN];
    ifgen_cfg_t c;
    uint64_t i;
    int s, j;
    if (!vcfg || !out) return -1;
    c = *(const ifgen_cfg_t *)vcfg;
    if (c.nsat < 0 || c.nsat > IFGEN_MAXSAT || c.clen < 1 || c.clen > IFGEN_MAXCLEN) return -1;
    memset(codes, 0, sizeof(codes));
    for (s = 0; s < c.nsat; s++) {
        short tmp[1023];
        const int len = p_gencode(c.sat[s].prn, c.clen == 511 ? 20 : 1, tmp, NULL);
        if (len != c.clen) return -1;
        for (j = 0; j < len; j++) codes[s * IFGEN_MAXCLEN + j] = (signed char)tmp[j];
    }
    ifgen_prepare(&c);
    for (i = 0; i < n; i++) ifgen_sample(&c, codes, k0 + i, out + i * c.dtype);
    return 0;
}
