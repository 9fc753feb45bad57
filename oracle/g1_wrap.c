/* g1_wrap.c - TEST INFRASTRUCTURE.  The reference's GLONASS C/A generator
 * gencode_G1G2() (src/sdrcode.c:426-444) is `static` and this fork's gencode()
 * switch (src/sdrcode.c:523-539) has no CTYPE_G1 case, so nothing in the
 * reference build can call it.  This translation unit textually includes the
 * reference source WHERE IT LIES (nothing is copied into the repository) so that
 * the static function becomes reachable; oracle/Makefile localises every other
 * symbol of the resulting object so that it does not clash with sdrcode.o.     */
#include "sdrcode.c"

int oh_gencode_g1_static(short *out, double *crate)
{
    int len = 0, i;
    double cr = 0.0;
    short *c = gencode_G1G2(&len, &cr);
    if (!c) return -1;
    for (i = 0; i < len; i++) out[i] = c[i];
    free(c);
    if (crate) *crate = cr;
    return len;
}
