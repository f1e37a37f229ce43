/* Exhaustive check of the division-by-constant shortcut of csrc/btf.cu (div_const): for every dividend the blur / mRTV
 * stage can produce, x * RN(1/c) refined by two FMAs equals the IEEE quotient x / c.  Built and run by tests/test_abi.py. */
#include <math.h>
#include <stdio.h>
static int check(float c, int xmax, long* n) {
    const float rc = 1.0f / c;
    int bad = 0;
    for (int x = 0; x <= xmax; ++x) {
        const float xf = (float)x, q0 = xf * rc, r = fmaf(-c, q0, xf), q = fmaf(r, rc, q0);
        if (q != xf / c) ++bad;
        ++*n;
    }
    return bad;
}
int main(void) {
    long n = 0;
    int bad = check(3.f, 765, &n);
    for (int K = 1; K <= 15; K += 2) bad += check((float)(K * K), K * K * 255, &n);
    printf("checked %ld mismatches %d\n", n, bad);
    return bad != 0;
}
