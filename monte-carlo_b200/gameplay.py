"""Mirror of the showdown part of montecarlo.gameplay (/root/reference/src/montecarlo/gameplay.clj:104-133).

`end_game_ranking` is the pure part of `end-game` (:127-132): value per remaining player, sorted descending,
partitioned into tie groups (sets of ids, best first) -- computed by one mc_showdown_groups call."""
from . import api
from .card import card_id


def group_winners(sorted_entries):                           # :118-120
    """[{player, hand_value}, ...] already sorted best-first -> list of sets of players, ties together"""
    groups, last = [], object()
    for e in sorted_entries:
        if groups and e["hand_value"] == last:
            groups[-1].add(e["player"])
        else:
            groups.append({e["player"]})
        last = e["hand_value"]
    return groups


def end_game_ranking(community_cards, hands):                # :127-132
    """community_cards: 3..5 deck Cards; hands: {player_id: [Card, Card]} -> (values {id: [cat hit kick]},
    winners = list of sets of ids, best group first)"""
    ids = list(hands.keys())
    board = api.mask_of_ids(card_id(c) for c in community_cards)
    holes = [api.mask_of_ids(card_id(c) for c in hands[p]) for p in ids]
    keys, _winners, group_of, n_groups = api.showdown(board, holes)
    values = {p: api.key_decode(k) for p, k in zip(ids, keys)}
    groups = [set() for _ in range(n_groups)]
    for p, g in zip(ids, group_of):
        groups[g].add(p)
    return values, groups


def update_stacks(pots, winners, stacks):                    # :104-116
    """pots: [{bet, n, players(set)}]; winners: tie groups best first; stacks: {id: chips} updated in place.
    The first tie group that intersects a pot's players splits int(total/|group|) each (:108-112) -- settled by
    mc_settle_pots.  A pot none of whose players is in any group raises (the reference divides by zero there, :112)."""
    ids = sorted({p for g in winners for p in g} | {p for pot in pots for p in pot["players"]}, key=repr)
    if len(ids) > 9:
        raise ValueError("more than nine players")
    seat = {p: i for i, p in enumerate(ids)}
    group_of = [-1] * len(ids)
    for gi, g in enumerate(winners):
        for p in g:
            group_of[seat[p]] = gi
    delta = api.settle_pots(group_of, [(pot["bet"], pot["n"], [seat[p] for p in pot["players"]]) for pot in pots]) if ids else []
    for p, d in zip(ids, delta):
        if d or p in stacks:
            stacks[p] = stacks.get(p, 0) + d
    return stacks
