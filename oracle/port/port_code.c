/* port_code.c - ranging codes in the reference's +-1 register arithmetic
 * (TEST INFRASTRUCTURE; see port.h).  C/A: src/sdrcode.c:101-154; GLONASS ST
 * code: src/sdrcode.c:426-444 (defined there but never dispatched,
 * src/sdrcode.c:523-539 - the port dispatches it so config 3 has an oracle). */
#include "port.h"

static const short k_delay[210] = {            /* G2 delay, src/sdrcode.c:104-126 */
    5,6,7,8,17,18,139,140,141,251,252,254,255,256,257,258,469,470,471,472,
    473,474,509,512,513,514,515,516,859,860,861,862,863,950,947,948,950,67,103,91,
    19,679,225,625,946,638,161,1001,554,280,710,709,775,864,558,220,397,55,898,759,
    367,299,1018,729,695,780,801,788,732,34,320,327,389,407,525,405,221,761,260,326,
    955,653,699,422,188,438,959,539,879,677,586,153,792,814,446,264,1015,278,536,819,
    156,957,159,712,885,461,248,713,126,807,279,122,197,693,632,771,467,647,203,145,
    175,52,21,237,235,886,657,634,762,355,1012,176,603,130,359,595,68,386,797,456,
    499,883,307,127,211,121,118,163,628,853,484,289,811,202,1021,463,568,904,670,230,
    911,684,309,644,932,12,314,891,212,185,675,503,150,395,345,846,798,992,357,995,
    877,112,144,476,193,109,445,291,87,399,292,901,339,208,711,189,263,537,663,942,
    173,900,30,500,935,556,373,85,652,310
};

static int gen_l1ca(int prn, short *code)
{
    signed char g1[1023], g2[1023], r1[10], r2[10];
    int i, j;
    if (prn < 1 || prn > 210) return -1;
    for (i = 0; i < 10; i++) r1[i] = r2[i] = -1;
    for (i = 0; i < 1023; i++) {
        const signed char f1 = (signed char)(r1[2] * r1[9]);
        const signed char f2 = (signed char)(r2[1] * r2[2] * r2[5] * r2[7] * r2[8] * r2[9]);
        g1[i] = r1[9]; g2[i] = r2[9];
        for (j = 9; j > 0; j--) { r1[j] = r1[j - 1]; r2[j] = r2[j - 1]; }
        r1[0] = f1; r2[0] = f2;
    }
    for (i = 0; i < 1023; i++) code[i] = (short)(-g1[i] * g2[(i + 1023 - k_delay[prn - 1]) % 1023]);
    return 1023;
}

static int gen_g1(short *code)
{
    signed char r[9];
    int i, j;
    for (i = 0; i < 9; i++) r[i] = -1;
    for (i = 0; i < 511; i++) {
        const signed char f = (signed char)(r[4] * r[8]);
        code[i] = (short)(-r[6]);
        for (j = 8; j > 0; j--) r[j] = r[j - 1];
        r[0] = f;
    }
    return 511;
}

int p_gencode(int prn, int ctype, short *code, double *crate)
{
    if (ctype == 1 || ctype == 27) { if (crate) *crate = 1.023e6; return gen_l1ca(prn, code); }
    if (ctype == 20) { if (crate) *crate = 0.511e6; return gen_g1(code); }
    return -1;
}
