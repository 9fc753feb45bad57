/* vendor_stubs.c - TEST INFRASTRUCTURE.  The reference's front-end driver files
 * (src/rcv/stereo/stereo.c, src/rcv/rtlsdr/rtlsdr.c, src/rcv/gn3s/gn3s.cpp) are
 * compiled WHERE THEY LIE for the sake of their byte-unpack routines
 * (stereo_exp, rtlsdr_exp, gn3s_exp_v3), which are plain table look-ups.  Their
 * other functions talk to USB devices through vendor libraries that do not exist
 * here (libnslstereo, librtlsdr, the GN3S FX2 class); those entry points are
 * satisfied by the do-nothing definitions below and are never called.           */
#include <stddef.h>

/* libnslstereo */
unsigned int STEREO_DATABUFF_SIZE = 0;
unsigned char *STEREO_dataBuffer = NULL;
int STEREO_GrabClean(void) { return -1; }
int STEREO_GrabStop(void) { return -1; }
int STEREO_InitLibrary(void) { return -1; }
int STEREO_IsConnected(void) { return 0; }
int STEREO_QuitLibrary(void) { return -1; }

/* librtlsdr + convenience.c */
int rtlsdr_open(void **dev, unsigned int index) { (void)dev; (void)index; return -1; }
int rtlsdr_close(void *dev) { (void)dev; return -1; }
int rtlsdr_cancel_async(void *dev) { (void)dev; return -1; }
int rtlsdr_read_async(void *dev, void *cb, void *ctx, unsigned int n, unsigned int len)
{ (void)dev; (void)cb; (void)ctx; (void)n; (void)len; return -1; }
int verbose_auto_gain(void *dev) { (void)dev; return -1; }
int verbose_device_search(char *s) { (void)s; return -1; }
int verbose_ppm_set(void *dev, int ppm) { (void)dev; (void)ppm; return -1; }
int verbose_reset_buffer(void *dev) { (void)dev; return -1; }
int verbose_set_frequency(void *dev, unsigned int f) { (void)dev; (void)f; return -1; }
int verbose_set_sample_rate(void *dev, unsigned int r) { (void)dev; (void)r; return -1; }

/* GN3S FX2 device class (C++ member functions, by their mangled names) */
#define STUB(name, mangled) int name(void) __asm__(mangled); int name(void) { return -1; }
STUB(fx2_xfer2, "_ZN7Fx2_dev10usrp_xfer2EhhPhc")
STUB(fx2_find, "_ZN7Fx2_dev12usb_fx2_findEv")
STUB(fx2_init, "_ZN7Fx2_dev12usb_fx2_initEv")
STUB(fx2_overrun, "_ZN7Fx2_dev16check_rx_overrunEPb")
STUB(fx2_close, "_ZN7Fx2_dev5closeEv")
STUB(fx2_read, "_ZN7Fx2_dev7read_IFEPh")
STUB(fx2_xfer, "_ZN7Fx2_dev9usrp_xferEhh")
STUB(fx2_ctor_i, "_ZN7Fx2_devC1Ei")
STUB(fx2_ctor, "_ZN7Fx2_devC1Ev")
STUB(fx2_dtor, "_ZN7Fx2_devD1Ev")
