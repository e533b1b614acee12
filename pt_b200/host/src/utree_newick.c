/*
 * pt_b200/host/src/utree_newick.c -- the host build's unrooted-Newick reader.
 *
 * The reference's reader is a bison/flex pair (LIBPLL/parse_utree.y,
 * LIBPLL/lex_utree.l); a build without those tools needs the six entry points
 * that live in the bison file from somewhere, so they are written here by hand
 * as a recursive-descent reader over libpll's public pll_unode_t / pll_utree_t
 * types (compiled against libpll's pll.h; host-only pointer work, no likelihood
 * code).  The ptw drivers of this repository link it; so does the in-place
 * build of the reference that serves as test oracle (oracle/Makefile.ref).
 *
 * Behaviour followed (citations into /root/reference/lib/libptpll/lib/
 * pll-modules/libs/libpll/src/):
 *   grammar               parse_utree.y:177-248  root = 3 subtrees, inner = 2,
 *                         optional label + optional ":number", final ';'
 *   tokens                lex_utree.l:80-122
 *   index assignment      parse_utree.y:250-340
 *   node array order      parse_utree.y:342-359,395-445
 *   destroy               parse_utree.y:37-137
 */
#include <libpll/pll.h>

typedef enum { T_END, T_OPAR, T_CPAR, T_COMMA, T_COLON, T_SEMI,
               T_STRING, T_NUMBER, T_BAD } tok_kind;

typedef struct {
  const char *p;        /* cursor */
  tok_kind kind;        /* look-ahead token */
  char *text;           /* malloc'd text for STRING/NUMBER */
  unsigned int tips;
  int failed;
} reader;

static int is_break(int c)
{
  return c == 0 || c == ' ' || c == '\t' || c == '\n' || c == '\r' ||
         c == '(' || c == ')' || c == '[' || c == ']' || c == ',' ||
         c == ':' || c == ';';
}

/* does s[0..n) match  [+-]?(d+ .? d* | d* .? d+)([eE][+-]?d+)?  or [+-]?d+ */
static int looks_numeric(const char *s, size_t n)
{
  size_t i = 0, nd = 0;
  if (i < n && (s[i] == '+' || s[i] == '-')) ++i;
  while (i < n && isdigit((unsigned char)s[i])) { ++i; ++nd; }
  if (i < n && s[i] == '.') ++i;
  while (i < n && isdigit((unsigned char)s[i])) { ++i; ++nd; }
  if (!nd) return 0;
  if (i < n && (s[i] == 'e' || s[i] == 'E'))
  {
    size_t j = i + 1, ne = 0;
    if (j < n && (s[j] == '+' || s[j] == '-')) ++j;
    while (j < n && isdigit((unsigned char)s[j])) { ++j; ++ne; }
    if (!ne) return 0;
    i = j;
  }
  return i == n;
}

static char *dup_range(const char *s, size_t n)
{
  char *m = (char *)malloc(n + 1);
  memcpy(m, s, n);
  m[n] = 0;
  return m;
}

static void advance(reader *r)
{
  r->text = NULL;
  for (;;)
  {
    int c = (unsigned char)*r->p;
    if (c == ' ' || c == '\t' || c == '\r' || c == '\n') { ++r->p; continue; }
    break;
  }
  int c = (unsigned char)*r->p;
  switch (c)
  {
    case 0:   r->kind = T_END;   return;
    case '(': r->kind = T_OPAR;  ++r->p; return;
    case ')': r->kind = T_CPAR;  ++r->p; return;
    case ',': r->kind = T_COMMA; ++r->p; return;
    case ':': r->kind = T_COLON; ++r->p; return;
    case ';': r->kind = T_SEMI;  ++r->p; return;
    case '[': case ']': r->kind = T_BAD; return;
    default: break;
  }
  if (c == '\'' || c == '"')
  {
    /* quoted label: runs to the matching quote; a backslash keeps the next
       character (lex_utree.l:80-99) */
    const char q = (char)c;
    const char *s = ++r->p;
    while (*r->p && *r->p != q)
    {
      if (*r->p == '\\' && r->p[1]) ++r->p;
      ++r->p;
    }
    if (!*r->p) { r->kind = T_BAD; return; }
    r->text = dup_range(s, (size_t)(r->p - s));
    ++r->p;
    r->kind = T_STRING;
    return;
  }
  const char *s = r->p;
  while (!is_break((unsigned char)*r->p)) ++r->p;
  size_t n = (size_t)(r->p - s);
  r->text = dup_range(s, n);
  r->kind = looks_numeric(s, n) ? T_NUMBER : T_STRING;
}

static pll_unode_t *new_node(void)
{
  return (pll_unode_t *)calloc(1, sizeof(pll_unode_t));
}

/* optional_label optional_length ; returns 0 on syntax error */
static int read_label_length(reader *r, char **label, double *length)
{
  *label = NULL;
  *length = 0;
  if (r->kind == T_STRING || r->kind == T_NUMBER)
  {
    *label = r->text;
    advance(r);
  }
  if (r->kind == T_COLON)
  {
    advance(r);
    if (r->kind != T_NUMBER) { free(r->text); return 0; }
    *length = atof(r->text);
    free(r->text);
    advance(r);
  }
  return 1;
}

static void free_partial(pll_unode_t *n)
{
  if (!n) return;
  if (n->next)
  {
    if (n->next->back) free_partial(n->next->back);
    if (n->next->next && n->next->next->back) free_partial(n->next->next->back);
    free(n->next->next);
    free(n->next);
  }
  free(n->label);
  free(n);
}

static pll_unode_t *read_subtree(reader *r)
{
  if (r->kind == T_OPAR)
  {
    advance(r);
    pll_unode_t *a = read_subtree(r);
    if (!a) return NULL;
    if (r->kind != T_COMMA) { free_partial(a); return NULL; }
    advance(r);
    pll_unode_t *b = read_subtree(r);
    if (!b) { free_partial(a); return NULL; }
    if (r->kind != T_CPAR) { free_partial(a); free_partial(b); return NULL; }
    advance(r);

    pll_unode_t *n = new_node();
    n->next = new_node();
    n->next->next = new_node();
    n->next->next->next = n;
    n->next->back = a;        a->back = n->next;
    n->next->next->back = b;  b->back = n->next->next;
    char *lab; double len;
    if (!read_label_length(r, &lab, &len)) { free_partial(n); return NULL; }
    n->label = n->next->label = n->next->next->label = lab;
    n->length = len;
    n->next->length = a->length;
    n->next->next->length = b->length;
    return n;
  }
  if (r->kind == T_STRING || r->kind == T_NUMBER)
  {
    pll_unode_t *n = new_node();
    char *lab; double len;
    if (!read_label_length(r, &lab, &len)) { free(n); return NULL; }
    n->label = lab;
    n->length = len;
    r->tips++;
    return n;
  }
  return NULL;
}

static pll_unode_t *read_tree(const char *s, unsigned int *tips)
{
  reader r;
  memset(&r, 0, sizeof r);
  r.p = s;
  advance(&r);
  if (r.kind != T_OPAR) { free(r.text); return NULL; }
  advance(&r);
  pll_unode_t *sub[3] = { NULL, NULL, NULL };
  int ok = 1;
  for (int i = 0; i < 3 && ok; ++i)
  {
    sub[i] = read_subtree(&r);
    if (!sub[i]) { ok = 0; break; }
    tok_kind want = (i < 2) ? T_COMMA : T_CPAR;
    if (r.kind != want) { ok = 0; break; }
    advance(&r);
  }
  char *lab = NULL; double len = 0;
  if (ok) ok = read_label_length(&r, &lab, &len);
  if (ok && r.kind != T_SEMI) ok = 0;
  if (ok) { advance(&r); if (r.kind != T_END) ok = 0; }
  if (!ok)
  {
    free(r.text);
    free(lab);
    for (int i = 0; i < 3; ++i) free_partial(sub[i]);
    pll_errno = PLL_ERROR_NEWICK_SYNTAX;
    snprintf(pll_errmsg, 200, "Syntax error in Newick string");
    return NULL;
  }
  pll_unode_t *root = new_node();
  root->next = new_node();
  root->next->next = new_node();
  root->next->next->next = root;
  root->back = sub[0];             sub[0]->back = root;
  root->next->back = sub[1];       sub[1]->back = root->next;
  root->next->next->back = sub[2]; sub[2]->back = root->next->next;
  root->label = root->next->label = root->next->next->label = lab;
  root->length = sub[0]->length;
  root->next->length = sub[1]->length;
  root->next->next->length = sub[2]->length;
  *tips = r.tips;
  return root;
}

/* ---- index template (parse_utree.y:250-340) ---- */
typedef struct { unsigned int tip, inner_clv, inner_node; int scaler; } counters;

static void number_subtree(pll_unode_t *n, counters *c)
{
  if (!n->next)
  {
    n->node_index = n->clv_index = n->pmatrix_index = c->tip++;
    n->scaler_index = PLL_SCALE_BUFFER_NONE;
    return;
  }
  number_subtree(n->next->back, c);
  number_subtree(n->next->next->back, c);
  pll_unode_t *ring[3] = { n, n->next, n->next->next };
  for (int k = 0; k < 3; ++k)
  {
    ring[k]->node_index = c->inner_node + (unsigned int)k;
    ring[k]->clv_index = c->inner_clv;
    ring[k]->scaler_index = c->scaler;
  }
  n->pmatrix_index = c->inner_clv;
  ring[1]->pmatrix_index = ring[1]->back->pmatrix_index;
  ring[2]->pmatrix_index = ring[2]->back->pmatrix_index;
  c->inner_clv++; c->scaler++; c->inner_node += 3;
}

PLL_EXPORT void pll_utree_reset_template_indices(pll_unode_t *node,
                                                 unsigned int tip_count)
{
  counters c = { 0, tip_count, tip_count, 0 };
  pll_unode_t *ring[3] = { node, node->next, node->next->next };
  for (int k = 0; k < 3; ++k) number_subtree(ring[k]->back, &c);
  for (int k = 0; k < 3; ++k)
  {
    ring[k]->node_index = c.inner_node + (unsigned int)k;
    ring[k]->clv_index = c.inner_clv;
    ring[k]->scaler_index = c.scaler;
    ring[k]->pmatrix_index = ring[k]->back->pmatrix_index;
  }
}

/* ---- node array (parse_utree.y:342-359,395-445) ---- */
static unsigned int count_tips(pll_unode_t *n)
{
  if (!n->next) return 1;
  return count_tips(n->next->back) + count_tips(n->next->next->back);
}

static void collect(pll_unode_t *n, pll_unode_t **arr,
                    unsigned int *ti, unsigned int *ii)
{
  if (!n->next) { arr[(*ti)++] = n; return; }
  collect(n->next->back, arr, ti, ii);
  collect(n->next->next->back, arr, ti, ii);
  arr[(*ii)++] = n;
}

PLL_EXPORT pll_utree_t *pll_utree_wraptree(pll_unode_t *root,
                                           unsigned int tip_count)
{
  if (tip_count < 3 && tip_count != 0)
  {
    pll_errno = PLL_ERROR_PARAM_INVALID;
    snprintf(pll_errmsg, 200, "Invalid tip_count value (%u).", tip_count);
    return NULL;
  }
  if (tip_count == 0)
  {
    if (!root->next && !root->back->next)
    {
      pll_errno = PLL_ERROR_PARAM_INVALID;
      snprintf(pll_errmsg, 200, "Input tree contains no inner nodes.");
      return NULL;
    }
    if (!root->next) root = root->back;
    tip_count = count_tips(root->back) + count_tips(root->next->back) +
                count_tips(root->next->next->back);
  }
  pll_utree_t *t = (pll_utree_t *)malloc(sizeof(pll_utree_t));
  t->nodes = (pll_unode_t **)malloc((2 * tip_count - 2) * sizeof(pll_unode_t *));
  unsigned int ti = 0, ii = tip_count;
  collect(root->back, t->nodes, &ti, &ii);
  collect(root->next->back, t->nodes, &ti, &ii);
  collect(root->next->next->back, t->nodes, &ti, &ii);
  t->nodes[ii] = root;
  t->tip_count = tip_count;
  t->inner_count = tip_count - 2;
  t->edge_count = 2 * tip_count - 3;
  return t;
}

PLL_EXPORT pll_utree_t *pll_utree_parse_newick_string(const char *s)
{
  unsigned int tips = 0;
  pll_unode_t *root = read_tree(s, &tips);
  if (!root) return NULL;
  pll_utree_reset_template_indices(root, tips);
  return pll_utree_wraptree(root, tips);
}

PLL_EXPORT pll_utree_t *pll_utree_parse_newick(const char *filename)
{
  FILE *f = fopen(filename, "r");
  if (!f)
  {
    pll_errno = PLL_ERROR_FILE_OPEN;
    snprintf(pll_errmsg, 200, "Unable to open file (%s)", filename);
    return NULL;
  }
  fseek(f, 0, SEEK_END);
  long n = ftell(f);
  fseek(f, 0, SEEK_SET);
  char *buf = (char *)malloc((size_t)n + 1);
  size_t got = fread(buf, 1, (size_t)n, f);
  buf[got] = 0;
  fclose(f);
  pll_utree_t *t = pll_utree_parse_newick_string(buf);
  free(buf);
  return t;
}

/* ---- destroy (parse_utree.y:37-137) ---- */
static void drop_data(pll_unode_t *n, void (*cb)(void *))
{
  if (n->data && cb) cb(n->data);
}

static void drop_subtree(pll_unode_t *n, void (*cb)(void *))
{
  if (n->next)
  {
    drop_subtree(n->next->back, cb);
    drop_subtree(n->next->next->back, cb);
    drop_data(n->next, cb);
    drop_data(n->next->next, cb);
    free(n->next->next);
    free(n->next);
  }
  drop_data(n, cb);
  free(n->label);
  free(n);
}

PLL_EXPORT void pll_utree_graph_destroy(pll_unode_t *root, void (*cb)(void *))
{
  if (!root) return;
  if (root->next)
  {
    drop_subtree(root->next->back, cb);
    drop_subtree(root->next->next->back, cb);
  }
  if (root->back) drop_subtree(root->back, cb);
  if (root->next)
  {
    drop_data(root->next, cb);
    drop_data(root->next->next, cb);
    free(root->next->next);
    free(root->next);
  }
  drop_data(root, cb);
  free(root->label);
  free(root);
}

PLL_EXPORT void pll_utree_destroy(pll_utree_t *tree, void (*cb)(void *))
{
  unsigned int i;
  for (i = 0; i < tree->tip_count + tree->inner_count; ++i)
  {
    pll_unode_t *n = tree->nodes[i];
    if (n->next)
    {
      drop_data(n->next, cb);
      drop_data(n->next->next, cb);
      free(n->next->next);
      free(n->next);
    }
    drop_data(n, cb);
    free(n->label);
    free(n);
  }
  free(tree->nodes);
  free(tree);
}
