/* TEST INFRASTRUCTURE (oracle) — included twice by azb_oracle.c (REAL = float / double).
 *
 * Cell-list neighbour-list builder emitting the HOOMD-blue v5 arrays the force kernels consume (SURVEY.md
 * Appendix A.6): n_neigh[i] (u32), head_list[i] (size_t row offset), nlist[head+k] (u32 particle INDEX, may be a
 * ghost >= N). FULL storage mode (i->j and j->i). A pair is listed when r < r_list[ti*T+tj] (= r_cut + r_buff,
 * <= 0 disables the type pair) and j is not in i's exclusion row. Rows are sorted by ascending j so that the
 * result is deterministic; HOOMD's own row order is cell-traversal order, parity is per-row SET equality.
 */

typedef struct
    {
    unsigned int key, val;
    } SUF(cellent);

static int SUF(cmp_u32)(const void* a, const void* b)
    {
    const unsigned int x = *(const unsigned int*)a, y = *(const unsigned int*)b;
    return (x > y) - (x < y);
    }

static int SUF(nlist_build)(const void* pos_v, unsigned int N, unsigned int n_max, const azb_box* box, const double* r_list,
                            unsigned int T, const size_t* excl_head, const unsigned int* excl_list, unsigned int row_align,
                            unsigned int* n_neigh, size_t* head_list, unsigned int** out_nlist, size_t* out_size)
    {
    const SUF(real4)* pos = (const SUF(real4)*)pos_v;
    double rmax = 0;
    for (unsigned int t = 0; t < T * T; ++t)
        if (r_list[t] > rmax)
            rmax = r_list[t];
    REAL* rlsq = (REAL*)malloc(sizeof(REAL) * T * T);
    for (unsigned int t = 0; t < T * T; ++t)
        rlsq[t] = r_list[t] > 0 ? (REAL)r_list[t] * (REAL)r_list[t] : (REAL)-1;
    if (row_align == 0)
        row_align = 1;

    const int brute = (box->xy != 0 || box->xz != 0 || box->yz != 0 || rmax <= 0);
    double lo[3], ext[3];
    int nc[3] = {1, 1, 1};
    const double L[3] = {box->Lx, box->Ly, box->Lz};
    for (int d = 0; d < 3; ++d)
        {
        if (box->periodic[d])
            {
            lo[d] = -0.5 * L[d];
            ext[d] = L[d];
            }
        else
            {
            double mn = 1e300, mx = -1e300;
            for (unsigned int i = 0; i < n_max; ++i)
                {
                const double c = d == 0 ? pos[i].x : (d == 1 ? pos[i].y : pos[i].z);
                if (c < mn)
                    mn = c;
                if (c > mx)
                    mx = c;
                }
            if (n_max == 0)
                mn = mx = 0;
            lo[d] = mn;
            ext[d] = (mx - mn) * (1.0 + 1e-9) + 1e-9;
            }
        if (!brute)
            {
            nc[d] = (int)floor(ext[d] / rmax);
            if (nc[d] < 1)
                nc[d] = 1;
            if (nc[d] > 1024)
                nc[d] = 1024;
            }
        }
    const size_t ncell = (size_t)nc[0] * nc[1] * nc[2];
    unsigned int* cell_of = (unsigned int*)malloc(sizeof(unsigned int) * (n_max ? n_max : 1));
    size_t* cell_start = (size_t*)calloc(ncell + 1, sizeof(size_t));
    unsigned int* cell_part = (unsigned int*)malloc(sizeof(unsigned int) * (n_max ? n_max : 1));
    for (unsigned int i = 0; i < n_max; ++i)
        {
        int c[3];
        const double p[3] = {pos[i].x, pos[i].y, pos[i].z};
        for (int d = 0; d < 3; ++d)
            {
            c[d] = (int)floor((p[d] - lo[d]) / ext[d] * nc[d]);
            if (c[d] < 0)
                c[d] = 0;
            if (c[d] >= nc[d])
                c[d] = nc[d] - 1;
            }
        cell_of[i] = (unsigned int)(((size_t)c[2] * nc[1] + c[1]) * nc[0] + c[0]);
        cell_start[cell_of[i] + 1]++;
        }
    for (size_t c = 0; c < ncell; ++c)
        cell_start[c + 1] += cell_start[c];
    size_t* fill = (size_t*)malloc(sizeof(size_t) * ncell);
    memcpy(fill, cell_start, sizeof(size_t) * ncell);
    for (unsigned int i = 0; i < n_max; ++i)
        cell_part[fill[cell_of[i]]++] = i;
    free(fill);

    unsigned int* nl = NULL;
    size_t total = 0;
    for (int pass = 0; pass < 2; ++pass)
        {
        if (pass == 1)
            {
            total = 0;
            for (unsigned int i = 0; i < N; ++i)
                {
                head_list[i] = total;
                total += ((size_t)n_neigh[i] + row_align - 1) / row_align * row_align;
                }
            nl = (unsigned int*)malloc(sizeof(unsigned int) * (total ? total : 1));
            memset(nl, 0, sizeof(unsigned int) * (total ? total : 1));
            }
#pragma omp parallel for schedule(dynamic, 512)
        for (unsigned int i = 0; i < N; ++i)
            {
            const unsigned int ti = (unsigned int)SUF(type_of)(pos[i].w);
            unsigned int cnt = 0;
            unsigned int* row = pass ? nl + head_list[i] : NULL;
            int ci[3];
                {
                unsigned int c = cell_of[i];
                ci[0] = (int)(c % nc[0]);
                c /= nc[0];
                ci[1] = (int)(c % nc[1]);
                ci[2] = (int)(c / nc[1]);
                }
            /* neighbouring cells, de-duplicated when a periodic dimension has fewer than 3 cells */
            int lst[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, nl_d[3];
            for (int d = 0; d < 3; ++d)
                {
                nl_d[d] = 0;
                for (int o = -1; o <= 1; ++o)
                    {
                    int c = ci[d] + o;
                    if (box->periodic[d])
                        c = (c % nc[d] + nc[d]) % nc[d];
                    else if (c < 0 || c >= nc[d])
                        continue;
                    int dup = 0;
                    for (int q = 0; q < nl_d[d]; ++q)
                        if (lst[d][q] == c)
                            dup = 1;
                    if (!dup)
                        lst[d][nl_d[d]++] = c;
                    }
                }
            for (int az = 0; az < nl_d[2]; ++az)
                for (int ay = 0; ay < nl_d[1]; ++ay)
                    for (int ax = 0; ax < nl_d[0]; ++ax)
                        {
                        const size_t c = ((size_t)lst[2][az] * nc[1] + lst[1][ay]) * nc[0] + lst[0][ax];
                        for (size_t s = cell_start[c]; s < cell_start[c + 1]; ++s)
                            {
                            const unsigned int j = cell_part[s];
                            if (j == i)
                                continue;
                            const unsigned int tj = (unsigned int)SUF(type_of)(pos[j].w);
                            const REAL rl = rlsq[ti * T + tj];
                            if (rl < 0)
                                continue;
                            REAL dx = pos[i].x - pos[j].x, dy = pos[i].y - pos[j].y, dz = pos[i].z - pos[j].z;
                            SUF(min_image)(box, &dx, &dy, &dz);
                            if (dx * dx + dy * dy + dz * dz >= rl)
                                continue;
                            if (excl_head)
                                {
                                int ex = 0;
                                for (size_t e = excl_head[i]; e < excl_head[i + 1]; ++e)
                                    if (excl_list[e] == j)
                                        ex = 1;
                                if (ex)
                                    continue;
                                }
                            if (pass)
                                row[cnt] = j;
                            ++cnt;
                            }
                        }
            if (pass)
                qsort(row, cnt, sizeof(unsigned int), SUF(cmp_u32));
            else
                n_neigh[i] = cnt;
            }
        }
    free(rlsq);
    free(cell_of);
    free(cell_start);
    free(cell_part);
    *out_nlist = nl;
    *out_size = total;
    return 0;
    }
